"""Parity of the tcgen05 implicit-GEMM convolution (through the C ABI) against a plain PyTorch fp32 conv2d.

Inputs and weights are bf16-representable, so the only differences are fp32 accumulation order and the final
rounding of the output to bf16: tolerance = 1 bf16 ulp of the result (2^-8 relative) + a small absolute term.
"""
import pytest
import torch
import torch.nn.functional as F

import rdb_b200
from rdb_b200.lib import check, ptr, stream_ptr

pytestmark = pytest.mark.gpu

ACT = {"none": 0, "relu": 1, "silu": 2}


def _case_seed(case):
    """Stable across processes (Python's hash() of a tuple holding strings is randomised per process)."""
    import zlib
    return zlib.crc32(repr(case).encode()) % (2 ** 31)


def ref_conv(x, w, b, res, stride, pad, act):
    y = F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), b, stride=stride, padding=pad)
    if res is not None:
        y = y + res.float().permute(0, 3, 1, 2)
    if act == "relu":
        y = F.relu(y)
    elif act == "silu":
        y = F.silu(y)
    return y.permute(0, 2, 3, 1).contiguous()


def run_conv(lib, x, w, b, res, stride, pad, act):
    n, h, wd, cin = x.shape
    cout, k = w.shape[0], w.shape[1]
    ho, wo = (h + 2 * pad - k) // stride + 1, (wd + 2 * pad - k) // stride + 1
    out = torch.full((n, ho, wo, cout), float("nan"), device=x.device, dtype=torch.bfloat16)
    check(lib.b200_conv2d_nhwc(ptr(x), n, h, wd, cin, 0, ptr(w), ptr(b), ptr(res), 0, ptr(out), cout, 0, k, stride,
                               pad, ACT[act], stream_ptr()), "b200_conv2d_nhwc")
    torch.cuda.synchronize()
    return out


CASES = [
    # n, h, w, cin, cout, k, stride, pad, act, residual
    (2, 8, 8, 64, 64, 1, 1, 0, "none", False),       # one k-block, one tile
    (2, 56, 56, 64, 256, 1, 1, 0, "relu", True),     # ResNet layer1 conv3 + residual
    (3, 28, 28, 512, 128, 1, 1, 0, "relu", False),   # multi k-block, ragged M (2352 = 18.4 tiles)
    (2, 56, 56, 64, 64, 3, 1, 1, "relu", False),     # 3x3, 8x8x2 tiles
    (4, 14, 14, 256, 256, 3, 1, 1, "relu", False),   # 3x3 at 14x14
    (3, 7, 7, 512, 512, 3, 1, 1, "relu", False),     # 3x3 at 7x7 (padded tiles, odd batch)
    (2, 56, 56, 128, 128, 3, 2, 1, "relu", False),   # stride-2 3x3 (ResNet v1.5 layer2.0.conv2)
    (2, 56, 56, 256, 512, 1, 2, 0, "none", False),   # stride-2 1x1 downsample
    (1, 28, 28, 24, 96, 3, 1, 1, "silu", False),     # EfficientNet-style odd channels (Cin not multiple of 64)
    (2, 14, 14, 160, 200, 1, 1, 0, "silu", True),    # Cout not a multiple of 64, residual
    (1, 15, 15, 64, 64, 3, 2, 1, "silu", False),     # odd spatial size with stride 2
    (2, 14, 14, 256, 1024, 1, 1, 0, "relu", True),   # CTA-pair kernel (N=256) with residual (layer3 conv3)
    (1, 7, 7, 512, 512, 1, 1, 0, "relu", False),     # CTA-pair kernel with a single (odd) m-tile: phantom peer tile
    (5, 14, 14, 1024, 256, 1, 1, 0, "relu", False),  # CTA-pair, 16 k-blocks, odd number of m-tiles (980 px -> 8 tiles) 
    (3, 28, 28, 128, 128, 3, 1, 1, "silu", True),    # CTA-pair N=128 with post-activation residual geometry (3x3)
    (2, 7, 7, 512, 2048, 1, 1, 0, "relu", True),     # CTA pair + residual, K = 512 (layer4 conv3): two epilogue warpgroups
    (3, 64, 64, 64, 64, 3, 1, 1, "silu", True),      # halo-tile 3x3 kernel with residual, H a multiple of 16
    (1, 16, 24, 64, 64, 3, 1, 1, "none", False),     # halo-tile kernel, one tile high, three wide
    (5, 56, 56, 64, 64, 3, 1, 1, "relu", False),     # halo-tile kernel, odd batch (tile count not a multiple of anything)
    (2, 20, 20, 64, 64, 3, 1, 1, "relu", False),     # W not a multiple of 8 / H waste too high: generic kernel
    (2, 112, 112, 24, 24, 3, 1, 1, "silu", True),    # halo kernel, 64-byte pixels (24 -> 32 channels zero-filled by TMA), N = 32
    (1, 56, 56, 48, 192, 3, 1, 1, "silu", False),    # halo kernel, 128-byte pixels, three resident channel tiles
    (2, 160, 160, 16, 16, 3, 1, 1, "silu", False),   # halo kernel, 32-byte pixels (YOLOv8n c2f bottleneck at 160x160)
    (3, 80, 80, 32, 32, 3, 1, 1, "silu", True),      # halo kernel, 64-byte pixels, residual
    (2, 30, 30, 8, 40, 3, 1, 1, "relu", False),      # halo kernel, Cin = 8, W not a multiple of 8, Cout 40 -> N = 64
    (5, 48, 40, 48, 136, 3, 1, 1, "none", True),     # halo kernel, odd batch, three channel tiles, last one ragged (136 = 2*64 + 8)
    (1, 17, 9, 16, 200, 3, 1, 1, "relu", False),     # too much tile padding for the halo kernel: generic kernel, thin K
    (2, 320, 320, 16, 32, 3, 2, 1, "silu", False),   # halo kernel, stride 2 through parity planes (YOLOv8n layer 1), 32-byte pixels
    (3, 160, 160, 32, 64, 3, 2, 1, "silu", False),   # halo kernel, stride 2, 64-byte pixels (YOLOv8n layer 3)
    (2, 112, 112, 24, 96, 3, 2, 1, "silu", False),   # halo kernel, stride 2, two channel tiles (EfficientNetV2 features.2.0)
    (1, 31, 45, 8, 24, 3, 2, 1, "relu", False),      # halo kernel, stride 2, odd input sizes (ragged parity planes)
    (2, 64, 48, 32, 40, 3, 2, 1, "none", True),      # halo kernel, stride 2 with a residual
]


@pytest.mark.parametrize("case", CASES, ids=[str(c) for c in CASES])
def test_conv_matches_fp32_reference(b200lib, cuda, case):
    n, h, w, cin, cout, k, stride, pad, act, use_res = case
    g = torch.Generator(device="cpu").manual_seed(_case_seed(case))
    x = torch.randn(n, h, w, cin, generator=g).to(cuda).to(torch.bfloat16)
    wt = (torch.randn(cout, k, k, cin, generator=g) / (k * k * cin) ** 0.5).to(cuda).to(torch.bfloat16)
    b = torch.randn(cout, generator=g).to(cuda)
    ho, wo = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
    res = torch.randn(n, ho, wo, cout, generator=g).to(cuda).to(torch.bfloat16) if use_res else None
    out = run_conv(b200lib, x, wt, b, res, stride, pad, act)
    ref = ref_conv(x, wt, b, res, stride, pad, act)
    assert not torch.isnan(out.float()).any(), "unwritten output elements"
    err = (out.float() - ref).abs()
    tol = ref.abs() * 2.0 ** -8 + 2e-3
    bad = (err > tol).sum().item()
    assert bad == 0, "max err %.4g (tol %.4g), %d bad of %d" % (err.max().item(), tol.max().item(), bad, err.numel())


@pytest.mark.parametrize("case", [(2, 56, 56, 64, 64, 1, 256, "relu"),      # layer1.0: conv3 + downsample, both stride 1, flat
                                  (3, 28, 28, 128, 256, 2, 512, "relu"),    # layer2.0: downsample reads every second pixel
                                  (2, 7, 7, 512, 1024, 2, 2048, "relu"),    # layer4.0: CTA pairs, 24 k-blocks
                                  (1, 14, 14, 64, 40, 2, 72, "none"),       # ragged second input (40 ch), odd sizes below
                                  (2, 7, 9, 128, 24, 2, 136, "silu")])
def test_conv1x1_dual_input(b200lib, cuda, case):
    """out = act(W[:, :ca] . a + W[:, ca:] . b[::s, ::s] + bias) as one K-concatenated GEMM vs two fp32 convs."""
    n, h, w, ca, cb, s2, cout, act = case
    g = torch.Generator(device="cpu").manual_seed(_case_seed(case))
    h2, w2 = (h * s2 - (s2 - 1), w * s2 - (s2 - 1)) if case[1] % 2 else (h * s2, w * s2)
    a = torch.randn(n, h, w, ca, generator=g).to(cuda).to(torch.bfloat16)
    b2 = torch.randn(n, h2, w2, cb, generator=g).to(cuda).to(torch.bfloat16)
    wt = (torch.randn(cout, ca + cb, generator=g) / (ca + cb) ** 0.5).to(cuda).to(torch.bfloat16)
    bias = torch.randn(cout, generator=g).to(cuda)
    out = torch.full((n, h, w, cout), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_conv1x1_dual_nhwc(ptr(a), n, h, w, ca, 0, ptr(b2), h2, w2, cb, 0, s2, ptr(wt), ptr(bias), ptr(out),
                                         cout, 0, ACT[act], stream_ptr()), "b200_conv1x1_dual_nhwc")
    torch.cuda.synchronize()
    wf = wt.float()
    y = F.conv2d(a.float().permute(0, 3, 1, 2), wf[:, :ca, None, None], bias) + \
        F.conv2d(b2.float().permute(0, 3, 1, 2), wf[:, ca:, None, None], None, stride=s2)
    y = {"none": lambda v: v, "relu": F.relu, "silu": F.silu}[act](y).permute(0, 2, 3, 1)
    assert not torch.isnan(out.float()).any(), "unwritten output elements"
    err = (out.float() - y).abs()
    tol = y.abs() * 2.0 ** -8 + 2e-3
    assert (err > tol).sum().item() == 0, "max err %.4g" % err.max().item()


@pytest.mark.parametrize("case", [(128, 14, 14, 960, 160, True), (5, 7, 7, 1536, 256, True), (3, 14, 14, 512, 128, False),
                                  (2, 9, 5, 200, 72, False)])
def test_conv1x1_with_folded_se_scale(b200lib, cuda, case):
    """out = W . bf16(x * gate[n][c]) + bias (+ residual): identical to channel_scale followed by the plain conv."""
    n, h, w, cin, cout, use_res = case
    g = torch.Generator(device="cpu").manual_seed(_case_seed(case))
    x = torch.randn(n, h, w, cin, generator=g).to(cuda).to(torch.bfloat16)
    gate = torch.sigmoid(torch.randn(n, cin, generator=g)).to(cuda)
    wt = (torch.randn(cout, 1, 1, cin, generator=g) / cin ** 0.5).to(cuda).to(torch.bfloat16)
    bias = torch.randn(cout, generator=g).to(cuda)
    res = torch.randn(n, h, w, cout, generator=g).to(cuda).to(torch.bfloat16) if use_res else None
    out = torch.full((n, h, w, cout), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_conv1x1_scaled_nhwc(ptr(x), n, h, w, cin, 0, ptr(gate), ptr(wt), ptr(bias), ptr(res), 0, ptr(out), cout,
                                           0, ACT["none"], stream_ptr()), "b200_conv1x1_scaled_nhwc")
    scaled = torch.empty_like(x)
    check(b200lib.b200_channel_scale_nhwc(ptr(x), n, h, w, cin, 0, ptr(gate), ptr(scaled), 0, stream_ptr()))
    two_pass = run_conv(b200lib, scaled, wt, bias, res, 1, 0, "none")
    torch.cuda.synchronize()
    assert not torch.isnan(out.float()).any(), "unwritten output elements"
    assert torch.equal(out.view(torch.int16), two_pass.view(torch.int16)), (out.float() - two_pass.float()).abs().max().item()
    ref = ref_conv(scaled, wt, bias, res, 1, 0, "none")
    err = (out.float() - ref).abs()
    assert (err > ref.abs() * 2.0 ** -8 + 2e-3).sum().item() == 0


def test_conv1x1_dual_rejects_bad_shapes(b200lib, cuda):
    a = torch.zeros(1, 8, 8, 32, device=cuda, dtype=torch.bfloat16)      # first input not a multiple of 64 channels
    b2 = torch.zeros(1, 8, 8, 64, device=cuda, dtype=torch.bfloat16)
    wt = torch.zeros(64, 96, device=cuda, dtype=torch.bfloat16)
    bias = torch.zeros(64, device=cuda)
    out = torch.empty(1, 8, 8, 64, device=cuda, dtype=torch.bfloat16)
    assert b200lib.b200_conv1x1_dual_nhwc(ptr(a), 1, 8, 8, 32, 0, ptr(b2), 8, 8, 64, 0, 1, ptr(wt), ptr(bias), ptr(out), 64, 0,
                                          0, stream_ptr()) != 0
    a = torch.zeros(1, 8, 8, 64, device=cuda, dtype=torch.bfloat16)
    assert b200lib.b200_conv1x1_dual_nhwc(ptr(a), 1, 8, 8, 64, 0, ptr(b2), 9, 9, 64, 0, 1, ptr(wt), ptr(bias), ptr(out), 64, 0,
                                          0, stream_ptr()) != 0          # geometry mismatch


def test_conv_channel_slices(b200lib, cuda):
    """Reads a channel slice of a wider buffer and writes into a slice of another (chunk / concat without copies)."""
    g = torch.Generator(device="cpu").manual_seed(7)
    n, h, w = 2, 14, 14
    xin = torch.randn(n, h, w, 128, generator=g).to(cuda).to(torch.bfloat16)
    wt = (torch.randn(64, 1, 1, 64, generator=g) / 8).to(cuda).to(torch.bfloat16)
    b = torch.randn(64, generator=g).to(cuda)
    out = torch.zeros(n, h, w, 192, device=cuda, dtype=torch.bfloat16)
    x_slice = xin[..., 64:128]
    o_slice = out[..., 128:192]
    check(b200lib.b200_conv2d_nhwc(x_slice.data_ptr(), n, h, w, 64, 128, ptr(wt), ptr(b), None, 0,
                                   o_slice.data_ptr(), 64, 192, 1, 1, 0, 1, stream_ptr()))
    torch.cuda.synchronize()
    ref = ref_conv(x_slice.contiguous(), wt, b, None, 1, 0, "relu")
    assert (out[..., :128] == 0).all()
    err = (out[..., 128:].float() - ref).abs()
    assert (err <= ref.abs() * 2.0 ** -8 + 2e-3).all(), err.max().item()


def test_conv_rejects_bad_shapes(b200lib, cuda):
    x = torch.zeros(1, 8, 8, 12, device=cuda, dtype=torch.bfloat16)
    rc = b200lib.b200_conv2d_nhwc(ptr(x), 1, 8, 8, 12, 0, ptr(x), ptr(x), None, 0, ptr(x), 12, 0, 1, 1, 0, 0, None)
    assert rc == -1
    assert b"multiples of 8" in b200lib.b200_last_error()


GATHER_CASES = [
    # n, h, w, cin, cout, k, stride, pad, act, residual, res_after_act
    (2, 112, 112, 24, 24, 3, 1, 1, "silu", True, True),     # EfficientNetV2 features.1 FusedMBConv (expand 1)
    (2, 112, 112, 24, 96, 3, 2, 1, "silu", False, False),   # features.2.0 expand3x3, stride 2
    (3, 56, 56, 48, 192, 3, 1, 1, "silu", False, False),    # features.2.x expand3x3
    (1, 160, 160, 16, 16, 3, 1, 1, "silu", True, True),     # YOLOv8n C2f bottleneck, c = 16
    (2, 80, 80, 32, 32, 3, 1, 1, "silu", False, False),     # YOLOv8n c = 32
    (1, 33, 47, 16, 32, 3, 2, 1, "relu", False, False),     # ragged sizes, stride 2
    (5, 9, 7, 56, 200, 3, 1, 1, "none", True, False),       # widest supported cin, odd cout
]


@pytest.mark.parametrize("case", GATHER_CASES, ids=[str(c) for c in GATHER_CASES])
def test_gather_conv_matches_fp32_reference(b200lib, cuda, case):
    n, h, w, cin, cout, k, stride, pad, act, use_res, after = case
    g = torch.Generator(device="cpu").manual_seed(_case_seed(case))
    x = torch.randn(n, h, w, cin, generator=g).to(cuda).to(torch.bfloat16)
    wt = (torch.randn(cout, k, k, cin, generator=g) / (k * k * cin) ** 0.5).to(torch.bfloat16)
    b = torch.randn(cout, generator=g)
    npad = b200lib.b200_gather_conv_npad(cout)
    kpad = b200lib.b200_gather_conv_kpad(cin, k)
    wp = torch.zeros(npad, kpad, dtype=torch.bfloat16)
    wp[:cout, :k * k * cin] = wt.reshape(cout, k * k * cin)
    bp = torch.zeros(npad)
    bp[:cout] = b
    wp, bp, wt, b = wp.to(cuda), bp.to(cuda), wt.to(cuda), b.to(cuda)
    ho, wo = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
    res = torch.randn(n, ho, wo, cout, generator=g).to(cuda).to(torch.bfloat16) if use_res else None
    out = torch.full((n, ho, wo, cout), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_gather_conv2d_nhwc(ptr(x), n, h, w, cin, 0, ptr(wp), ptr(bp), ptr(res), 0, ptr(out), cout, 0, k,
                                          stride, pad, ACT[act], 1 if after else 0, stream_ptr()), "b200_gather_conv2d_nhwc")
    torch.cuda.synchronize()
    y = F.conv2d(x.float().permute(0, 3, 1, 2), wt.float().permute(0, 3, 1, 2), b, stride=stride, padding=pad)
    fa = {"none": lambda t: t, "relu": F.relu, "silu": F.silu}[act]
    if res is not None:
        r = res.float().permute(0, 3, 1, 2)
        y = fa(y) + r if after else fa(y + r)
    else:
        y = fa(y)
    ref = y.permute(0, 2, 3, 1)
    assert not torch.isnan(out.float()).any(), "unwritten output elements"
    err = (out.float() - ref).abs()
    tol = ref.abs() * 2.0 ** -8 + 2e-3
    assert (err <= tol).all(), "max err %.4g" % err.max().item()


SHUFFLE_CASES = [
    # n, h, w, cin (physical), c, cp
    (3, 28, 28, 64, 58, 64),        # stage 2: both half-groups in one 64-slot tile (two 32-channel boxes of the other branch)
    (5, 14, 14, 128, 116, 128),     # stage 3: one tile per half-group, ragged M (980 pixels)
    (2, 7, 7, 256, 232, 256),       # stage 4: two tiles per half-group, a single partial M tile
    (64, 28, 28, 64, 58, 64),       # 392 tiles on 148 CTAs: the staging ring wraps
    (128, 14, 14, 128, 116, 128),   # the same with two channel tiles
    (1, 5, 5, 8, 40, 64),           # thin K, 20 valid slots per half-group
]


@pytest.mark.parametrize("case", SHUFFLE_CASES, ids=[str(c) for c in SHUFFLE_CASES])
def test_conv1x1_shuffle_tail(b200lib, cuda, case):
    """channel_shuffle(cat(other, relu(conv1x1(x))), 2) from the conv's own epilogue (tv shufflenetv2.py:24-38,96-103):
    the other branch's channels are copied bit for bit, the conv's meet the fp32 reference, pads are zero, and the whole
    tensor equals b200_conv2d_nhwc followed by b200_shuffle_concat_nhwc."""
    from torchvision.models.shufflenetv2 import channel_shuffle
    n, h, w, cin, c, cp = case
    g = torch.Generator(device="cpu").manual_seed(_case_seed(case))
    x = torch.randn(n, h, w, cin, generator=g).to(cuda).to(torch.bfloat16)
    # the other branch is the first half of the unit's input [x1 | x2] (pixel stride 2 cp); its pad channels hold junk
    parent = torch.randn(n, h, w, 2 * cp, generator=g).to(cuda).to(torch.bfloat16)
    wt = (torch.randn(c, cin, generator=g) / cin ** 0.5).to(torch.bfloat16)
    bias = torch.randn(c, generator=g)
    slot = torch.cat([torch.arange(c // 2), cp // 2 + torch.arange(c // 2)])
    wslots = torch.zeros(cp, cin, dtype=torch.bfloat16)
    wslots[slot] = wt
    bslots = torch.zeros(cp)
    bslots[slot] = bias
    wslots, bslots = wslots.to(cuda), bslots.to(cuda)
    out = torch.full((n, h, w, 2 * cp), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_conv1x1_shuffle_nhwc(ptr(x), n, h, w, cin, 0, ptr(parent), 2 * cp, c, cp, ptr(wslots), ptr(bslots),
                                            ptr(out), 0, stream_ptr()), "b200_conv1x1_shuffle_nhwc")
    torch.cuda.synchronize()
    assert not torch.isnan(out.float()).any(), "unwritten output elements"
    other = parent[..., :c]
    y = F.relu(F.conv2d(x.float().permute(0, 3, 1, 2), wt.to(cuda).float()[:, :, None, None], bias.to(cuda)))
    ref = channel_shuffle(torch.cat([other.float().permute(0, 3, 1, 2), y], dim=1), 2).permute(0, 2, 3, 1)
    got = torch.cat([out[..., :c], out[..., cp:cp + c]], dim=-1).float()
    assert torch.equal(got[..., 0::2], ref[..., 0::2]), "other branch not copied exactly"
    err = (got[..., 1::2] - ref[..., 1::2]).abs()
    tol = ref[..., 1::2].abs() * 2.0 ** -8 + 2e-3
    assert (err > tol).sum().item() == 0, "max err %.4g" % err.max().item()
    assert not out[..., c:cp].float().abs().any() and not out[..., cp + c:].float().abs().any(), "pads not zero"
    # the two-kernel path it replaces
    wplain = torch.zeros(cp, 1, 1, cin, dtype=torch.bfloat16, device=cuda)
    wplain[:c, 0, 0] = wt.to(cuda)
    bplain = torch.zeros(cp, device=cuda)
    bplain[:c] = bias.to(cuda)
    t = run_conv(b200lib, x, wplain, bplain, None, 1, 0, "relu")
    two_cp = (c + 7) // 8 * 8      # b200_shuffle_concat_nhwc pads its halves to 8 channels
    two = torch.full((n, h, w, 2 * two_cp), float("nan"), device=cuda, dtype=torch.bfloat16)
    a = parent[..., :cp].contiguous()
    check(b200lib.b200_shuffle_concat_nhwc(ptr(a), cp, ptr(t), cp, n, h, w, c, ptr(two), 0, stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(out[..., :c], two[..., :c]) and torch.equal(out[..., cp:cp + c], two[..., two_cp:two_cp + c])
