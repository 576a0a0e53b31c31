"""Operator-level parity of the bandwidth-bound kernels (through the C ABI) against plain PyTorch fp32 on the same
bf16-representable inputs.  Pure data movement and max-pooling are bit-exact; arithmetic kernels are compared within
1 bf16 ulp of the result (2^-8 relative + small absolute term), fp32 outputs within 1e-4 relative."""
import pytest
import torch
import torch.nn.functional as F

import rdb_b200
from rdb_b200.lib import check, ptr, stream_ptr

pytestmark = pytest.mark.gpu


def rnd(shape, seed, dev, dtype=torch.bfloat16):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(*shape, generator=g).to(dev).to(dtype)


def close_bf16(out, ref):
    err = (out.float() - ref).abs()
    return bool((err <= ref.abs() * 2.0 ** -8 + 2e-3).all()), err.max().item()


@pytest.mark.parametrize("k,s,p,cout,cfg", [(7, 2, 3, 64, (2, 224, 224)), (3, 2, 1, 24, (3, 224, 224)),
                                            (3, 2, 1, 16, (1, 97, 131)), (3, 2, 1, 32, (2, 64, 64))])
def test_stem_conv(b200lib, cuda, k, s, p, cout, cfg):
    n, h, w = cfg
    x4 = torch.zeros(n, h, w, 4, device=cuda, dtype=torch.bfloat16)
    x4[..., :3] = rnd((n, h, w, 3), 1, cuda)
    wt = rnd((cout, 3, k, k), 2, cuda, torch.float32) / (27 ** 0.5)
    b = rnd((cout,), 3, cuda, torch.float32)
    ct = b200lib.b200_stem_cout_tile(cout)
    cpad = (cout + ct - 1) // ct * ct
    wp = torch.zeros(k * k, 3, cpad, device=cuda)
    wp[:, :, :cout] = wt.permute(2, 3, 1, 0).reshape(k * k, 3, cout)
    bp = torch.zeros(cpad, device=cuda)
    bp[:cout] = b
    ho, wo = (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1
    out = torch.full((n, ho, wo, cout), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_stem_conv_nhwc(ptr(x4), n, h, w, ptr(wp), ptr(bp), ptr(out), cout, 0, k, s, p, 2, stream_ptr()))
    torch.cuda.synchronize()
    ref = F.silu(F.conv2d(x4[..., :3].float().permute(0, 3, 1, 2), wt, b, stride=s, padding=p)).permute(0, 2, 3, 1)
    ok, e = close_bf16(out, ref)
    assert ok, e


@pytest.mark.parametrize("cfg", [(2, 224, 224), (3, 64, 96), (1, 38, 304), (2, 640, 640)])
def test_stem_row_toeplitz(b200lib, cuda, cfg):
    """7x7/2 stem read in place through non-swizzled descriptors; includes widths that need several 128-pixel segments."""
    from rdb_b200.lowering import stem_row_weights
    n, h, w = cfg
    x4 = torch.zeros(n, h, w, 4, device=cuda, dtype=torch.bfloat16)
    x4[..., :3] = rnd((n, h, w, 3), 1, cuda)
    wt = (rnd((64, 3, 7, 7), 2, cuda, torch.float32) / 147 ** 0.5).to(torch.bfloat16).float()
    b = rnd((64,), 3, cuda, torch.float32)
    wb, bb = stem_row_weights(wt.cpu(), b.cpu())
    assert wb.nbytes == b200lib.b200_stem_row_weight_bytes()
    wdev = torch.from_numpy(wb.astype("int16")).to(cuda)
    bdev = torch.from_numpy(bb).to(cuda)
    ho, wo = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    out = torch.full((n, ho, wo, 64), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_stem_row_conv_nhwc(ptr(x4), n, h, w, ptr(wdev), ptr(bdev), ptr(out), 0, stream_ptr()))
    torch.cuda.synchronize()
    ref = F.relu(F.conv2d(x4[..., :3].float().permute(0, 3, 1, 2), wt, b, stride=2, padding=3)).permute(0, 2, 3, 1)
    assert not torch.isnan(out.float()).any()
    ok, e = close_bf16(out, ref)
    assert ok, e


@pytest.mark.parametrize("cfg", [(2, 224, 224), (3, 64, 96), (1, 40, 256), (5, 224, 224)])
def test_stem_pool_fused_bit_exact_vs_unfused_kernels(b200lib, cuda, cfg):
    """conv1+bn1+relu+maxpool in one kernel == Toeplitz stem kernel followed by the max-pool kernel, bit for bit."""
    from rdb_b200.lowering import stem_row_weights
    n, h, w = cfg
    x4 = torch.zeros(n, h, w, 4, device=cuda, dtype=torch.bfloat16)
    x4[..., :3] = rnd((n, h, w, 3), 1, cuda)
    wt = (rnd((64, 3, 7, 7), 2, cuda, torch.float32) / 147 ** 0.5).to(torch.bfloat16).float()
    b = rnd((64,), 3, cuda, torch.float32)
    wb, bb = stem_row_weights(wt.cpu(), b.cpu())
    wdev = torch.from_numpy(wb.astype("int16")).to(cuda)
    bdev = torch.from_numpy(bb).to(cuda)
    hs, ws = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    stem = torch.empty(n, hs, ws, 64, device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_stem_row_conv_nhwc(ptr(x4), n, h, w, ptr(wdev), ptr(bdev), ptr(stem), 0, stream_ptr()))
    pooled = torch.empty(n, hs // 2, ws // 2, 64, device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_maxpool_nhwc(ptr(stem), n, hs, ws, 64, 0, ptr(pooled), 0, 3, 2, 1, stream_ptr()))
    fused = torch.full((n, hs // 2, ws // 2, 64), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_stem_pool_nhwc(ptr(x4), n, h, w, ptr(wdev), ptr(bdev), ptr(fused), 0, stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(fused, pooled)
    ref = F.max_pool2d(F.relu(F.conv2d(x4[..., :3].float().permute(0, 3, 1, 2), wt, b, stride=2, padding=3)), 3, 2, 1)
    ok, e = close_bf16(fused, ref.permute(0, 2, 3, 1))
    assert ok, e


@pytest.mark.parametrize("k,cout,act,cfg", [(3, 24, 1, (3, 224, 224)), (3, 24, 2, (2, 224, 224)), (3, 16, 2, (1, 640, 640)),
                                            (3, 32, 2, (2, 64, 96)), (3, 48, 1, (1, 38, 304)), (7, 64, 1, (1, 64, 64))])
def test_stem_row_toeplitz_general(b200lib, cuda, k, cout, act, cfg):
    """The Toeplitz stem for every served stride-2 stem (3x3 pad 1 of ShuffleNetV2 / EfficientNetV2 / YOLOv8, 7x7 pad 3),
    plain and fused with MaxPool2d(3,2,1); channel counts below the accumulator width are clipped by the TMA store."""
    from rdb_b200.lowering import stem_row_weights
    n, h, w = cfg
    pad = k // 2
    x4 = torch.zeros(n, h, w, 4, device=cuda, dtype=torch.bfloat16)
    x4[..., :3] = rnd((n, h, w, 3), 1, cuda)
    wt = (rnd((cout, 3, k, k), 2, cuda, torch.float32) / (3 * k * k) ** 0.5).to(torch.bfloat16).float()
    b = rnd((cout,), 3, cuda, torch.float32)
    wb, bb = stem_row_weights(wt.cpu(), b.cpu())
    assert wb.nbytes == b200lib.b200_stem_row_weight_bytes_k(k, cout)
    wdev = torch.from_numpy(wb.astype("int16")).to(cuda)
    bdev = torch.from_numpy(bb).to(cuda)
    ho, wo = (h + 2 * pad - k) // 2 + 1, (w + 2 * pad - k) // 2 + 1
    out = torch.full((n, ho, wo, cout), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_stem_row_conv_k_nhwc(ptr(x4), n, h, w, ptr(wdev), ptr(bdev), ptr(out), cout, 0, k, act, stream_ptr()))
    torch.cuda.synchronize()
    ref = F.conv2d(x4[..., :3].float().permute(0, 3, 1, 2), wt, b, stride=2, padding=pad)
    ref = F.relu(ref) if act == 1 else (F.silu(ref) if act == 2 else ref)
    assert not torch.isnan(out.float()).any()
    ok, e = close_bf16(out, ref.permute(0, 2, 3, 1))
    assert ok, e
    if wo <= 128 and ho % 2 == 0 and wo % 2 == 0:
        pooled = torch.empty(n, ho // 2, wo // 2, cout, device=cuda, dtype=torch.bfloat16)
        check(b200lib.b200_maxpool_nhwc(ptr(out), n, ho, wo, cout, 0, ptr(pooled), 0, 3, 2, 1, stream_ptr()))
        fused = torch.full((n, ho // 2, wo // 2, cout), float("nan"), device=cuda, dtype=torch.bfloat16)
        check(b200lib.b200_stem_pool_k_nhwc(ptr(x4), n, h, w, ptr(wdev), ptr(bdev), ptr(fused), cout, 0, k, act,
                                            stream_ptr()))
        torch.cuda.synchronize()
        assert torch.equal(fused.view(torch.int16), pooled.view(torch.int16))


def test_stem_row_rejects_uninstantiated_combination(b200lib, cuda):
    x4 = torch.zeros(1, 32, 32, 4, device=cuda, dtype=torch.bfloat16)
    w = torch.zeros(b200lib.b200_stem_row_weight_bytes_k(7, 32) // 2, device=cuda, dtype=torch.int16)
    b = torch.zeros(64, device=cuda)
    out = torch.empty(1, 16, 16, 32, device=cuda, dtype=torch.bfloat16)
    assert b200lib.b200_stem_row_conv_k_nhwc(ptr(x4), 1, 32, 32, ptr(w), ptr(b), ptr(out), 32, 0, 7, 1, stream_ptr()) != 0
    assert b200lib.b200_stem_row_conv_k_nhwc(ptr(x4), 1, 32, 32, ptr(w), ptr(b), ptr(out), 32, 0, 5, 1, stream_ptr()) != 0


@pytest.mark.parametrize("k,p,cout,cfg,act", [(7, 3, 64, (2, 224, 224), 1), (3, 1, 24, (3, 224, 224), 2),
                                              (3, 1, 16, (1, 96, 130), 2), (7, 3, 64, (5, 38, 50), 1),
                                              (3, 1, 32, (2, 64, 64), 0)])
def test_stem_conv_tensor_core(b200lib, cuda, k, p, cout, cfg, act):
    from rdb_b200.lowering import stem_tc_weights
    n, h, w = cfg
    x4 = torch.zeros(n, h, w, 4, device=cuda, dtype=torch.bfloat16)
    x4[..., :3] = rnd((n, h, w, 3), 1, cuda)
    wt = (rnd((cout, 3, k, k), 2, cuda, torch.float32) / (3 * k * k) ** 0.5).to(torch.bfloat16).float()
    b = rnd((cout,), 3, cuda, torch.float32)
    wb, bb = stem_tc_weights(wt.cpu(), b.cpu())
    wdev = torch.from_numpy(wb.astype("int16")).to(cuda)
    bdev = torch.from_numpy(bb).to(cuda)
    ho, wo = (h + 2 * p - k) // 2 + 1, (w + 2 * p - k) // 2 + 1
    out = torch.full((n, ho, wo, cout), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_stem_conv_tc_nhwc(ptr(x4), n, h, w, ptr(wdev), ptr(bdev), ptr(out), cout, 0, k, 2, p, act,
                                         stream_ptr()))
    torch.cuda.synchronize()
    ref = F.conv2d(x4[..., :3].float().permute(0, 3, 1, 2), wt, b, stride=2, padding=p)
    ref = (F.relu(ref) if act == 1 else F.silu(ref) if act == 2 else ref).permute(0, 2, 3, 1)
    assert not torch.isnan(out.float()).any()
    ok, e = close_bf16(out, ref)
    assert ok, e


@pytest.mark.parametrize("n,h,w,c,stride,act", [(2, 28, 28, 64, 1, 0), (3, 56, 56, 24, 2, 0), (1, 15, 17, 960, 1, 2),
                                                (2, 14, 14, 232, 2, 2), (5, 14, 14, 232, 1, 1), (3, 7, 7, 464, 1, 0),
                                                (2, 28, 28, 120, 1, 2), (1, 9, 14, 40, 1, 1), (4, 42, 21, 72, 1, 0),
                                                # tile kernel: many items per block with the channel group changing inside a run,
                                                # ragged last group, two bands per image
                                                (160, 14, 14, 200, 1, 2), (300, 7, 7, 40, 1, 0), (41, 28, 28, 72, 1, 1),
                                                # stride-2 tiles: odd input sizes, 32-channel groups, several bands, ragged groups
                                                (3, 27, 28, 24, 2, 1), (70, 112, 112, 24, 2, 0), (9, 55, 56, 120, 2, 2),
                                                (40, 14, 14, 232, 2, 0), (2, 13, 13, 8, 2, 2)])
def test_dwconv3x3(b200lib, cuda, n, h, w, c, stride, act):
    x = rnd((n, h, w, c), 1, cuda)
    wt = rnd((c, 1, 3, 3), 2, cuda, torch.float32) / 3
    b = rnd((c,), 3, cuda, torch.float32)
    ho, wo = (h - 1) // stride + 1, (w - 1) // stride + 1
    out = torch.full((n, ho, wo, c), float("nan"), device=cuda, dtype=torch.bfloat16)
    w9 = wt.reshape(c, 9).t().contiguous()
    check(b200lib.b200_dwconv3x3_nhwc(ptr(x), n, h, w, c, 0, ptr(w9), ptr(b), ptr(out), 0, stride, act, stream_ptr()))
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt, b, stride=stride, padding=1, groups=c)
    ref = (F.silu(ref) if act == 2 else F.relu(ref) if act == 1 else ref).permute(0, 2, 3, 1)
    ok, e = close_bf16(out, ref)
    assert ok, e


@pytest.mark.parametrize("n,h,w,c,stride,act", [(3, 14, 14, 960, 1, 2), (2, 7, 7, 1536, 1, 2), (2, 28, 28, 256, 2, 2),
                                                (5, 14, 14, 960, 2, 2), (1, 21, 7, 8, 1, 0), (2, 13, 14, 16, 2, 1),
                                                (2, 28, 28, 24, 1, 2), (1, 56, 28, 40, 2, 0)])
def test_dwconv3x3_fused_squeeze_and_se_gate(b200lib, cuda, n, h, w, c, stride, act):
    """Strip kernel with the SE squeeze fused in (odd row counts, both strides), then the one-launch gate; against
    torch: AdaptiveAvgPool2d(1) of the STORED output, fc1 + SiLU, fc2 + sigmoid (tv ops/misc.py:249-254)."""
    x = rnd((n, h, w, c), 11, cuda)
    wt = rnd((c, 1, 3, 3), 12, cuda, torch.float32) / 3
    b = rnd((c,), 13, cuda, torch.float32)
    ho, wo = (h - 1) // stride + 1, (w - 1) // stride + 1
    strips = b200lib.b200_dwconv3x3_squeeze_rows(c, ho, wo, stride)
    raw = ((ho + 1) // 2 if stride == 1 else ho) * (wo // 7)
    assert strips in (raw, 1)      # 1: the tile kernel (or B200_DW_SQ=1) adds an image's strips itself
    out = torch.full((n, ho, wo, c), float("nan"), device=cuda, dtype=torch.bfloat16)
    sums = torch.full((n, strips, c), float("nan"), device=cuda)
    w9 = wt.reshape(c, 9).t().contiguous()
    check(b200lib.b200_dwconv3x3_squeeze_nhwc(ptr(x), n, h, w, c, 0, ptr(w9), ptr(b), ptr(out), 0, stride, act,
                                              ptr(sums), stream_ptr()))
    torch.cuda.synchronize()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt, b, stride=stride, padding=1, groups=c)
    ref = (F.silu(ref) if act == 2 else F.relu(ref) if act == 1 else ref).permute(0, 2, 3, 1)
    ok, e = close_bf16(out, ref)
    assert ok, e
    want_sum = out.float().sum(dim=(1, 2))
    assert torch.allclose(sums.sum(dim=1), want_sum, rtol=1e-5, atol=1e-4)
    # same output as the plain entry point, bit for bit
    out2 = torch.empty_like(out)
    check(b200lib.b200_dwconv3x3_nhwc(ptr(x), n, h, w, c, 0, ptr(w9), ptr(b), ptr(out2), 0, stride, act, stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(out.view(torch.int16), out2.view(torch.int16))

    sq = max(8, c // 24 // 8 * 8)
    w1 = rnd((sq, c), 14, cuda, torch.float32) / c ** 0.5
    b1 = rnd((sq,), 15, cuda, torch.float32)
    w2 = rnd((c, sq), 16, cuda, torch.float32) / sq ** 0.5
    b2 = rnd((c,), 17, cuda, torch.float32)
    w2t = w2.t().contiguous()
    gate = torch.full((n, c), float("nan"), device=cuda)
    check(b200lib.b200_se_gate(ptr(sums), n, strips, c, 1.0 / (ho * wo), ptr(w1), ptr(b1), ptr(w2t), ptr(b2), sq, ptr(gate),
                               stream_ptr()))
    torch.cuda.synchronize()
    mean = out.float().mean(dim=(1, 2))
    want = torch.sigmoid(F.silu(mean @ w1.t() + b1) @ w2.t() + b2)
    assert torch.allclose(gate, want, rtol=1e-4, atol=1e-5), (gate - want).abs().max().item()

    # the two-launch form the engine uses: hidden units, then fc2 + sigmoid + scale in one pass over the activation
    hid = torch.full((n, sq), float("nan"), device=cuda)
    check(b200lib.b200_se_hidden(ptr(sums), n, strips, c, 1.0 / (ho * wo), ptr(w1), ptr(b1), sq, ptr(hid), stream_ptr()))
    scaled = torch.full((n, ho, wo, c), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_se_scale_nhwc(ptr(out), n, ho, wo, c, 0, ptr(hid), ptr(w2t), ptr(b2), sq, ptr(scaled), 0,
                                     stream_ptr()))
    torch.cuda.synchronize()
    assert torch.allclose(hid, F.silu(mean @ w1.t() + b1), rtol=1e-4, atol=1e-5)
    ok, e = close_bf16(scaled, out.float() * want.view(n, 1, 1, c))
    assert ok, e


@pytest.mark.parametrize("n,c,sq,rows", [(19, 1536, 64, 1), (150, 960, 40, 1), (8, 512, 32, 3), (1, 256, 16, 1),
                                         (33, 40, 10, 2), (9, 1056, 44, 1), (3, 24, 8, 7), (5, 200, 13, 1)])
def test_se_gate_many_images_ragged_slices(b200lib, cuda, n, c, sq, rows):
    """The gate alone, at shapes the fused test does not reach: more image groups than clusters (150 images), channel
    counts that leave the last CTAs of a cluster a short or empty slice (40, 200, 1056), odd squeeze widths, several rows
    of sums per image.  Against torch in fp32 (tv ops/misc.py:249-254)."""
    g = torch.Generator().manual_seed(100 + n)
    sums = torch.randn(n, rows, c, generator=g).to(cuda)
    hw = 49
    w1 = (torch.randn(sq, c, generator=g) / c ** 0.5).to(cuda)
    b1 = torch.randn(sq, generator=g).to(cuda)
    w2 = (torch.randn(c, sq, generator=g) / sq ** 0.5).to(cuda)
    b2 = torch.randn(c, generator=g).to(cuda)
    w2t = w2.t().contiguous()
    gate = torch.full((n, c), float("nan"), device=cuda)
    check(b200lib.b200_se_gate(ptr(sums), n, rows, c, 1.0 / hw, ptr(w1), ptr(b1), ptr(w2t), ptr(b2), sq, ptr(gate),
                               stream_ptr()))
    torch.cuda.synchronize()
    mean = sums.sum(dim=1) / hw
    want = torch.sigmoid(F.silu(mean @ w1.t() + b1) @ w2.t() + b2)
    assert torch.isfinite(gate).all()
    assert torch.allclose(gate, want, rtol=1e-4, atol=2e-5), (gate - want).abs().max().item()
    # deterministic: a second launch gives the same bits
    gate2 = torch.empty_like(gate)
    check(b200lib.b200_se_gate(ptr(sums), n, rows, c, 1.0 / hw, ptr(w1), ptr(b1), ptr(w2t), ptr(b2), sq, ptr(gate2),
                               stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(gate, gate2)


def test_dwconv3x3_squeeze_rejects_widths_without_strips(b200lib, cuda):
    x = rnd((1, 10, 10, 8), 1, cuda)
    w9 = torch.zeros(9, 8, device=cuda)
    b = torch.zeros(8, device=cuda)
    out = torch.empty(1, 10, 10, 8, device=cuda, dtype=torch.bfloat16)
    sums = torch.empty(1, 16, 8, device=cuda)
    assert b200lib.b200_dwconv3x3_strips(10, 10, 1) == 0
    rc = b200lib.b200_dwconv3x3_squeeze_nhwc(ptr(x), 1, 10, 10, 8, 0, ptr(w9), ptr(b), ptr(out), 0, 1, 0, ptr(sums),
                                             stream_ptr())
    assert rc != 0


@pytest.mark.parametrize("n,h,w,c,k,s,p", [(2, 112, 112, 64, 3, 2, 1), (1, 20, 20, 128, 5, 1, 2), (3, 7, 9, 24, 3, 2, 1)])
def test_maxpool_bit_exact(b200lib, cuda, n, h, w, c, k, s, p):
    x = rnd((n, h, w, c), 4, cuda)
    ho, wo = (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1
    out = torch.empty(n, ho, wo, c, device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_maxpool_nhwc(ptr(x), n, h, w, c, 0, ptr(out), 0, k, s, p, stream_ptr()))
    torch.cuda.synchronize()
    ref = F.max_pool2d(x.float().permute(0, 3, 1, 2), k, s, p).permute(0, 2, 3, 1)
    assert torch.equal(out.float(), ref)


@pytest.mark.parametrize("n,h,w,c", [(5, 7, 7, 2048), (2, 7, 7, 1280), (3, 7, 7, 1024), (1, 14, 14, 96), (2, 3, 5, 8)])
def test_global_avgpool(b200lib, cuda, n, h, w, c):
    x = rnd((n, h, w, c), 5, cuda)
    out = torch.empty(n, c, device=cuda)
    check(b200lib.b200_global_avgpool_nhwc(ptr(x), n, h, w, c, 0, ptr(out), stream_ptr()))
    torch.cuda.synchronize()
    ref = x.float().mean(dim=(1, 2))
    assert torch.allclose(out, ref, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("n,k,j,act", [(1, 2048, 1000, 0), (19, 1280, 1000, 0), (256, 1024, 1000, 0), (7, 1536, 64, 2),
                                        (7, 40, 960, 3), (3, 16, 256, 3)])
def test_fc(b200lib, cuda, n, k, j, act):
    x = rnd((n, k), 6, cuda, torch.float32)
    w = rnd((j, k), 7, cuda, torch.float32) / k ** 0.5
    b = rnd((j,), 8, cuda, torch.float32)
    out = torch.full((n, j), float("nan"), device=cuda)
    check(b200lib.b200_fc(ptr(x), ptr(w), ptr(b), ptr(out), n, k, j, 1.0, act, stream_ptr()))
    torch.cuda.synchronize()
    ref = x.double() @ w.double().t() + b.double()
    ref = (F.silu(ref) if act == 2 else torch.sigmoid(ref) if act == 3 else ref).float()
    assert torch.allclose(out, ref, rtol=1e-4, atol=1e-5), (out - ref).abs().max().item()


def test_channel_scale(b200lib, cuda):
    n, h, w, c = 3, 14, 14, 960
    x = rnd((n, h, w, c), 9, cuda)
    gate = torch.sigmoid(rnd((n, c), 10, cuda, torch.float32))
    out = torch.empty_like(x)
    check(b200lib.b200_channel_scale_nhwc(ptr(x), n, h, w, c, 0, ptr(gate), ptr(out), 0, stream_ptr()))
    torch.cuda.synchronize()
    ref = (x.float() * gate.view(n, 1, 1, c)).to(torch.bfloat16)
    assert torch.equal(out, ref)


def test_copy_and_upsample_into_slices_bit_exact(b200lib, cuda):
    n, h, w, c = 2, 10, 12, 64
    x = rnd((n, h, w, c), 11, cuda)
    cat = torch.zeros(n, 2 * h, 2 * w, 192, device=cuda, dtype=torch.bfloat16)
    up_slice = cat[..., 128:192]
    check(b200lib.b200_upsample2x_nhwc(ptr(x), n, h, w, c, 0, up_slice.data_ptr(), 192, stream_ptr()))
    y = rnd((n, 2 * h, 2 * w, 128), 12, cuda)
    check(b200lib.b200_copy_channels_nhwc(ptr(y), n, 2 * h, 2 * w, 128, 0, cat.data_ptr(), 192, stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(cat[..., :128], y)
    assert torch.equal(cat[..., 128:], x.repeat_interleave(2, dim=1).repeat_interleave(2, dim=2))


@pytest.mark.parametrize("c", [58, 116, 232, 8])
def test_shuffle_concat_matches_torchvision_channel_shuffle(b200lib, cuda, c):
    from torchvision.models.shufflenetv2 import channel_shuffle
    n, h, w = 2, 7, 9
    cp = (c + 7) // 8 * 8
    a = torch.zeros(n, h, w, cp, device=cuda, dtype=torch.bfloat16)
    b = torch.zeros(n, h, w, cp, device=cuda, dtype=torch.bfloat16)
    a[..., :c] = rnd((n, h, w, c), 13, cuda)
    b[..., :c] = rnd((n, h, w, c), 14, cuda)
    out = torch.full((n, h, w, 2 * cp), float("nan"), device=cuda, dtype=torch.bfloat16)
    check(b200lib.b200_shuffle_concat_nhwc(ptr(a), 0, ptr(b), 0, n, h, w, c, ptr(out), 0, stream_ptr()))
    torch.cuda.synchronize()
    ref = channel_shuffle(torch.cat([a[..., :c], b[..., :c]], dim=-1).permute(0, 3, 1, 2).float(), 2).permute(0, 2, 3, 1)
    assert torch.equal(out[..., :c].float(), ref[..., :c])
    assert torch.equal(out[..., cp:cp + c].float(), ref[..., c:])
    assert not out[..., c:cp].float().abs().any() and not out[..., cp + c:].float().abs().any()
