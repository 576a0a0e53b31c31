"""Lowers torchvision-style eval-mode modules into the flat plan + weights blob that libb200infer executes.

What is fused here (so that every conv kernel reads its input once and writes its output once):
  * BatchNorm2d is folded into the preceding convolution (fp32 fold, then weights rounded to bf16, bias kept fp32);
  * ReLU / SiLU, the residual add of a block and channel concatenation (writing into a slice) go into the conv epilogue.

Blob layout (little endian): BlobHeader | BufRec[n_bufs] | OpRec[n_ops] | pad to 256 | weights section.
See csrc/engine.h for the record definitions; `parse_blob` below is the inverse (used by the CPU plan interpreter in
oracle/ to check the lowering without a GPU).
"""
import os
import struct
from dataclasses import dataclass
from typing import List, Optional

import numpy as np
import torch
import torch.nn as nn

OP_CONV, OP_STEM, OP_MAXPOOL, OP_AVGPOOL, OP_FC, OP_DWCONV, OP_SCALE, OP_COPY, OP_UPSAMPLE, OP_SHUFFLE, \
    OP_YOLO_DECODE, OP_GATHER_CONV, OP_SE_GATE, OP_SE_HIDDEN, OP_SE_SCALE = range(1, 16)
ACT_NONE, ACT_RELU, ACT_SILU, ACT_SIGMOID = 0, 1, 2, 3
HEADER_FMT = "<8sIIII32siiiiii3f3fQQ"
HEADER_SIZE = struct.calcsize(HEADER_FMT)
BUF_FMT = "<iiii"
OP_FMT = "<i23iqqqq"
assert struct.calcsize(OP_FMT) == 128

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def pad8(c):
    return (c + 7) // 8 * 8


def phys_channels(c):
    """Physical channel count of a fresh activation buffer. Multiples of 8 (16-byte chunks) everywhere, except that a
    24-channel tensor is stored with 32: its 48-byte pixels straddle 32-byte sectors and every TMA row read of it costs
    2-3 sectors for 1.5 (measured: 24-channel layers ran at 1.8 TB/s where 16- and 32-channel ones reach 4.5-4.7 TB/s).
    The extra channels are never written (buffers start zeroed) and meet zero weights. B200_PAD_24=0 turns it off."""
    if c == 24 and os.environ.get("B200_PAD_24", "1") != "0":
        return 32
    if c == 80 and os.environ.get("B200_PAD_80", "1") != "0":
        return 128      # YOLOv8n's class branch: 160-byte pixels -> 256-byte ones (measured: 1.262 -> 1.232 ms @32)
    return line_channels(c)


def line_channels(c):
    """pad8(c), or the next multiple of 64 channels (128-byte lines) when that costs at most 12 % more bytes: a TMA box row
    of 64 channels that starts at pixel * 240 bytes (120 channels) straddles two 128-byte lines on most pixels, and the
    thin 1x1 convs of ShuffleNetV2 (58 / 116 / 232-channel halves) were waiting on exactly those loads (ncu: the MMA warp
    and the producer both stalled on the A/B ring at 17 GB/s per SM). B200_PAD_LINE=0 turns it off."""
    p8 = pad8(c)
    p64 = (c + 63) // 64 * 64
    pct = int(os.environ.get("B200_PAD_LINE_PCT", "12"))
    if c > 32 and p64 != p8 and p64 * 100 <= p8 * (100 + pct) and os.environ.get("B200_PAD_LINE", "1") != "0":
        return p64
    return p8


def weight_rows(cout_p):
    """Rows a conv's weight matrix is stored with: the output channels rounded up to the channel tile the kernel will use
    (64, 128, or 256 for CTA pairs), so that no TMA box over the weights is partly out of bounds (see `weight_k`)."""
    if cout_p <= 64:
        return 64
    return (cout_p + 127) // 128 * 128 if cout_p < 256 else (cout_p + 255) // 256 * 256


def weight_k(cin_p, halo):
    """Input channels per filter tap the weights are stored with: rounded up to the K box of the kernel (64 channels; the
    halo 3x3 kernel uses 16 / 32 / 64). Measured on B200: a TMA box that is partly out of bounds loads about twice as slowly
    as one inside the tensor, and a thin layer loads its weight box once per 128-pixel tile. B200_PAD_WEIGHTS=0 stores the
    exact extents (the tensor map then zero-fills the rest of the box)."""
    if halo:
        return 16 if cin_p <= 16 else (32 if cin_p <= 32 else 64)
    return (cin_p + 63) // 64 * 64


def stem_cout_tile(cout):
    """Mirror of b200_stem_cout_tile (csrc/stem.cu)."""
    return 32 if cout % 32 == 0 else (24 if cout % 24 == 0 else 16)


@dataclass
class T:
    """A channel slice of an activation buffer: logical tensor [N, h, w, c] stored at channel offset `coff`.
    `cphys` physical channels are occupied (>= c, multiple of 8); logical channel l lives at physical offset cmap[l]
    inside the slice (identity unless the tensor is stored as padded groups, e.g. ShuffleNet's two halves)."""
    buf: int
    coff: int
    c: int
    h: int
    w: int
    cphys: int = 0
    cmap: Optional[np.ndarray] = None

    def __post_init__(self):
        if not self.cphys:
            self.cphys = pad8(self.c)
        if self.cmap is None:
            self.cmap = np.arange(self.c)


def stem_tc_weights(weight, bias):
    """[Cout,3,k,k] fp32 -> (bf16 bits [nout][kpad], fp32 bias [nout]) in the layout of csrc/stem_tc.cu:
    GEMM-K index = r * row_elems + (s + 1) * 4 + c, window pixel 0 / channel 3 / padded rows are zero."""
    cout, _, k, _ = weight.shape
    row_elems = 32 if k == 7 else 16
    kpad = 256 if k == 7 else 64
    nout = 16 if cout <= 16 else (32 if cout <= 32 else 64)
    w = torch.zeros(nout, kpad)
    for r in range(k):
        for s_ in range(k):
            base = r * row_elems + (s_ + 1) * 4
            w[:cout, base:base + 3] = weight[:, :, r, s_]
    b = torch.zeros(nout)
    b[:cout] = bias
    return PlanBuilder.bf16_bits(w), b.numpy().astype(np.float32)


def stem_row_weights(weight, bias):
    """[cout,3,k,k] fp32 (k = 7 or 3, cout <= 64) -> bf16 bits of the shared-memory image of csrc/stem_row.cu:
    [kernel row k][K chunk kc][nout][8] with kc = 4 (7x7) or 2 (3x3), nout = 32 or 64; element e of chunk c = window pixel
    2c + e // 4 (= kernel column + 1), channel e % 4. Bias padded to nout."""
    cout, cin, k, k2 = weight.shape
    assert cin == 3 and k == k2 and k in (3, 7) and cout <= 64
    kc = 4 if k == 7 else 2
    nout = 32 if cout <= 32 else 64
    img = torch.zeros(k, kc, nout, 8)
    for r in range(k):
        for s_ in range(k):
            wp = s_ + 1
            c, half = wp // 2, wp % 2
            img[r, c, :cout, half * 4:half * 4 + 3] = weight[:, :, r, s_]
    bp = torch.zeros(nout)
    bp[:cout] = bias
    return PlanBuilder.bf16_bits(img), bp.numpy().astype(np.float32)


def act_code(m):
    if m is None or isinstance(m, nn.Identity):
        return ACT_NONE
    if isinstance(m, nn.ReLU):
        return ACT_RELU
    if isinstance(m, nn.SiLU):
        return ACT_SILU
    if isinstance(m, nn.Sigmoid):
        return ACT_SIGMOID
    raise NotImplementedError("activation %r" % (m,))


def fold_bn(conv: nn.Conv2d, bn: Optional[nn.BatchNorm2d]):
    """fp32 fold of an eval-mode BatchNorm into the conv: returns (weight [Cout,Cin/g,kh,kw], bias [Cout]) in float64->32."""
    w = conv.weight.detach().double()
    b = conv.bias.detach().double() if conv.bias is not None else torch.zeros(w.shape[0], dtype=torch.float64)
    if bn is not None:
        scale = bn.weight.detach().double() / torch.sqrt(bn.running_var.detach().double() + bn.eps)
        w = w * scale.view(-1, 1, 1, 1)
        b = (b - bn.running_mean.detach().double()) * scale + bn.bias.detach().double()
    return w.float(), b.float()


class PlanBuilder:
    def __init__(self, kind, in_h, in_w, mean=IMAGENET_MEAN, std=IMAGENET_STD, pre_fill=0):
        self.kind = kind
        self.in_h, self.in_w = in_h, in_w
        self.mean, self.std, self.pre_fill = mean, std, pre_fill
        self.bufs = []      # (h, w, c, dtype)
        self.ops = []       # (type, p[23], w_off, b_off, w2_off, b2_off)
        self.weights = bytearray()
        self.in_buf = self.new_buf(in_h, in_w, 4)
        self.out_buf = -1
        self.out_dim = 0
        self.names = []     # op index -> human readable name (kept out of the blob)
        self.tensor_core_stem = True
        self.toeplitz_stem = True
        self.fuse_stem_pool = os.environ.get('B200_FUSE_STEM_POOL', '1') != '0'   # conv1 + bn1 + relu + maxpool in one kernel
        # ShuffleNetV2 cat + channel_shuffle written by the epilogue of the conv that ends branch2 (no shuffle pass)
        self.fuse_shuffle = os.environ.get('B200_FUSE_SHUFFLE', '1') != '0'
        self.pad_weights = os.environ.get('B200_PAD_WEIGHTS', '1') != '0'   # conv weights stored padded to the TMA box sizes
        self.gather_small_cin = os.environ.get('B200_GATHER_SMALL_CIN', '1') != '0'   # 3x3 convs with Cin <= 56 the halo kernel does not take: packed-K gather kernel
        self.fuse_downsample = os.environ.get('B200_FUSE_DOWNSAMPLE', '1') != '0'   # ResNet conv3 + downsample as one GEMM
        self.halo_small_cin = os.environ.get('B200_HALO_SMALL_CIN', '1') != '0'   # 3x3/s1 with Cin < 64 on the halo kernel
        self.halo_stride2 = os.environ.get('B200_CONV_HALO_S2', '1') != '0'      # 3x3/s2 with Cin <= 32 too (parity planes)
        self.gather_1x1_max_cin = int(__import__('os').environ.get('B200_GATHER_1X1_MAX_CIN', '0'))

    # ---- buffers / weights ----
    def new_buf(self, h, w, c, dtype=0):
        self.bufs.append((h, w, c, dtype))
        return len(self.bufs) - 1

    def new_act(self, h, w, c):
        cp = phys_channels(c)
        return T(self.new_buf(h, w, cp), 0, c, h, w, cp)

    def input(self):
        return T(self.in_buf, 0, 3, self.in_h, self.in_w)

    def add_weights(self, arr: np.ndarray):
        while len(self.weights) % 256:
            self.weights.append(0)
        off = len(self.weights)
        self.weights += np.ascontiguousarray(arr).tobytes()
        return off

    @staticmethod
    def bf16_bits(t: torch.Tensor) -> np.ndarray:
        return t.contiguous().to(torch.bfloat16).view(torch.int16).numpy().astype(np.uint16)

    def _op(self, typ, p, w_off=-1, b_off=-1, w2_off=-1, b2_off=-1, name=""):
        p = list(p) + [0] * (23 - len(p))
        self.ops.append((typ, p, w_off, b_off, w2_off, b2_off))
        self.names.append(name)

    # ---- ops ----
    def conv(self, x: T, weight, bias, stride, pad, act, res: Optional[T] = None, out: Optional[T] = None, name="",
             res_after_act=False, in_scale: Optional[T] = None):
        """weight: fp32 [Cout, Cin, k, k] (already BN-folded); x.c == Cin logical channels; physical channels padded to 8."""
        cout, cin, k, _ = weight.shape
        assert cin == x.c, (cin, x.c)
        ho, wo = (x.h + 2 * pad - k) // stride + 1, (x.w + 2 * pad - k) // stride + 1
        if out is None:
            out = self.new_act(ho, wo, cout)
        assert (out.h, out.w, out.c) == (ho, wo, cout)
        cin_p, cout_p = x.cphys, out.cphys
        halo = (self.halo_small_cin and k == 3 and pad == 1 and cout_p <= 256 and
                ((stride == 1 and cin_p <= 64) or (stride == 2 and self.halo_stride2 and cin_p <= 32 and x.w >= 2 and x.h >= 2))
                and ((wo + 7) // 8 * 8) * ((ho + 15) // 16 * 16) * 100 <= wo * ho * 135)   # conv_prepare's halo rule
        if in_scale is None and not halo and self.gather_small_cin and cout_p <= 256 and ((k == 3 and cin_p <= 56) or
                                                       (k == 1 and cin_p % 64 != 0 and cin_p <= self.gather_1x1_max_cin)):
            return self._gather_conv(x, weight, bias, stride, pad, act, res, out, name, res_after_act)
        kpad, rows = (weight_k(cin_p, halo), weight_rows(cout_p)) if self.pad_weights else (cin_p, cout_p)
        wp = torch.zeros(rows, k * k, kpad)
        wfull = torch.zeros(cout, k * k, kpad)
        wfull[:, :, torch.as_tensor(x.cmap)] = weight.permute(0, 2, 3, 1).reshape(cout, k * k, cin)
        wp[torch.as_tensor(out.cmap)] = wfull
        bp = torch.zeros(cout_p)
        bp[torch.as_tensor(out.cmap)] = bias
        w_off = self.add_weights(self.bf16_bits(wp))
        b_off = self.add_weights(bp.numpy().astype(np.float32))
        rb, rc = (res.buf, res.coff) if res is not None else (-1, 0)
        if res is not None:
            assert (res.h, res.w, res.c, res.cphys) == (ho, wo, cout, cout_p) and np.array_equal(res.cmap, out.cmap)
        pr = [x.buf, x.coff, cin_p, out.buf, out.coff, cout_p, rb, rc, k, stride, pad, act, 1 if res_after_act else 0]
        if in_scale is not None:      # SE gate [cin_p] fp32 folded into this conv's input (csrc/conv_gemm.cu, ConvSpec::a_scale)
            assert (k, stride, pad) == (1, 1, 0) and self.bufs[in_scale.buf][3] == 1 and in_scale.cphys == cin_p
            assert x.coff == 0 and self.bufs[x.buf][2] == cin_p
            pr += [0, 0, 0, 0, in_scale.buf + 1]
        if self.pad_weights:          # physical extent of the weights: [rows][k*k][kpad]
            pr += [0] * (19 - len(pr)) + [kpad, rows]
        self._op(OP_CONV, pr, w_off, b_off, name=name)
        return out

    def conv_dual(self, x: T, weight, bias, x2: T, weight2, bias2, stride2, act, name=""):
        """act(conv1x1(x) + conv1x1_stride2(x2)) as one K-concatenated GEMM (csrc/conv_gemm.cu, ConvSpec::in2): the last
        conv of a ResNet bottleneck and its downsample branch, whose output is then neither written nor re-read.
        weight [Cout, Ca, 1, 1], weight2 [Cout, Cb, 1, 1] (BN folded); biases are added."""
        cout, ca = weight.shape[:2]
        cb = weight2.shape[1]
        assert weight.shape[2:] == (1, 1) and weight2.shape[2:] == (1, 1) and weight2.shape[0] == cout
        assert ca == x.c and cb == x2.c and x.cphys % 64 == 0
        assert ((x2.h - 1) // stride2 + 1, (x2.w - 1) // stride2 + 1) == (x.h, x.w)
        out = self.new_act(x.h, x.w, cout)
        cap, cbp, cout_p = x.cphys, x2.cphys, out.cphys
        kpad, rows = (cap + weight_k(cbp, False), weight_rows(cout_p)) if self.pad_weights else (cap + cbp, cout_p)
        wp = torch.zeros(rows, kpad)
        wa = torch.zeros(cout, cap)
        wa[:, torch.as_tensor(x.cmap)] = weight.reshape(cout, ca)
        wb = torch.zeros(cout, cbp)
        wb[:, torch.as_tensor(x2.cmap)] = weight2.reshape(cout, cb)
        wp[torch.as_tensor(out.cmap), :cap + cbp] = torch.cat([wa, wb], dim=1)
        bp = torch.zeros(cout_p)
        bp[torch.as_tensor(out.cmap)] = bias + bias2
        w_off = self.add_weights(self.bf16_bits(wp))
        b_off = self.add_weights(bp.numpy().astype(np.float32))
        pr = [x.buf, x.coff, cap, out.buf, out.coff, cout_p, -1, 0, 1, 1, 0, act, 0, x2.buf + 1, x2.coff, cbp, stride2]
        if self.pad_weights:
            pr += [0] * (19 - len(pr)) + [kpad, rows]
        self._op(OP_CONV, pr, w_off, b_off, name=name)
        return out

    def _gather_conv(self, x, weight, bias, stride, pad, act, res, out, name, res_after_act):
        """Packed-K layout of csrc/gather_conv.cu: bf16 [npad][kpad], GEMM-K index = tap * cin_p + physical channel."""
        cout, cin, k, _ = weight.shape
        cin_p, cout_p = x.cphys, out.cphys
        npad = 32 if cout_p <= 32 else (64 if cout_p <= 64 else (128 if cout_p <= 128 else 256))
        kpad = (k * k * (cin_p // 8) + 7) // 8 * 64
        wfull = torch.zeros(cout, k * k, cin_p)
        wfull[:, :, torch.as_tensor(x.cmap)] = weight.permute(0, 2, 3, 1).reshape(cout, k * k, cin)
        wp = torch.zeros(npad, kpad)
        wp[torch.as_tensor(out.cmap), :k * k * cin_p] = wfull.reshape(cout, k * k * cin_p)
        bp = torch.zeros(npad)
        bp[torch.as_tensor(out.cmap)] = bias
        w_off = self.add_weights(self.bf16_bits(wp))
        b_off = self.add_weights(bp.numpy().astype(np.float32))
        rb, rc = (res.buf, res.coff) if res is not None else (-1, 0)
        if res is not None:
            assert (res.h, res.w, res.c, res.cphys) == (out.h, out.w, cout, cout_p) and np.array_equal(res.cmap, out.cmap)
        self._op(OP_GATHER_CONV, [x.buf, x.coff, cin_p, out.buf, out.coff, cout_p, rb, rc, k, stride, pad, act,
                                  1 if res_after_act else 0], w_off, b_off, name=name)
        return out

    def stem(self, x: T, weight, bias, stride, pad, act, name=""):
        cout, cin, k, _ = weight.shape
        assert cin == 3 and x.buf == self.in_buf
        ho, wo = (x.h + 2 * pad - k) // stride + 1, (x.w + 2 * pad - k) // stride + 1
        out = self.new_act(ho, wo, cout)
        if self.toeplitz_stem and self.toeplitz_ok(k, stride, pad, cout, act) and x.w % 16 == 0:
            wb, bb = stem_row_weights(weight, bias)
            w_off = self.add_weights(wb)
            b_off = self.add_weights(bb)
            self._op(OP_STEM, [x.buf, 0, 4, out.buf, out.coff, pad8(cout), -1, 0, k, stride, pad, act, 2], w_off, b_off,
                     name=name)
            return out
        if self.tensor_core_stem and stride == 2 and ((k == 7 and pad == 3) or (k == 3 and pad == 1)) and cout <= 64 \
                and x.w % 2 == 0:
            wb, bb = stem_tc_weights(weight, bias)
            w_off = self.add_weights(wb)
            b_off = self.add_weights(bb)
            self._op(OP_STEM, [x.buf, 0, 4, out.buf, out.coff, pad8(cout), -1, 0, k, stride, pad, act, 1], w_off, b_off,
                     name=name)
            return out
        ct = stem_cout_tile(pad8(cout))
        cpad = (pad8(cout) + ct - 1) // ct * ct
        wp = torch.zeros(k * k, 3, cpad)
        wp[:, :, :cout] = weight.permute(2, 3, 1, 0).reshape(k * k, 3, cout)
        bp = torch.zeros(cpad)
        bp[:cout] = bias
        w_off = self.add_weights(wp.numpy().astype(np.float32))
        b_off = self.add_weights(bp.numpy().astype(np.float32))
        self._op(OP_STEM, [x.buf, 0, 4, out.buf, out.coff, pad8(cout), -1, 0, k, stride, pad, act], w_off, b_off,
                 name=name)
        return out

    @staticmethod
    def toeplitz_ok(k, stride, pad, cout, act):
        """The (kernel, width, activation) combinations csrc/stem_row.cu instantiates (B200_STEM_DISPATCH)."""
        if stride != 2 or cout > 64:
            return False
        if (k, pad) == (7, 3):
            return cout > 32 and act == ACT_RELU
        if (k, pad) == (3, 1):
            return act in (ACT_RELU, ACT_SILU)
        return False

    def stem_pool(self, x: T, weight, bias, act=ACT_RELU, name=""):
        """stride-2 stem conv (7x7 pad 3 or 3x3 pad 1) + bias + act + MaxPool2d(3,2,1) in one kernel
        (csrc/stem_row.cu stem_pool_kernel): the stem activation never leaves shared memory."""
        cout, _, k, _ = weight.shape
        assert x.buf == self.in_buf and self.toeplitz_ok(k, 2, k // 2, cout, act)
        hs, ws = (x.h - 1) // 2 + 1, (x.w - 1) // 2 + 1
        assert ws <= 128 and hs % 2 == 0 and ws % 2 == 0
        out = self.new_act(hs // 2, ws // 2, cout)
        wb, bb = stem_row_weights(weight, bias)
        w_off = self.add_weights(wb)
        b_off = self.add_weights(bb)
        self._op(OP_STEM, [x.buf, 0, 4, out.buf, out.coff, pad8(cout), -1, 0, k, 2, k // 2, act, 3], w_off, b_off, name=name)
        return out

    @staticmethod
    def dw_strips(c, ho, wo, stride):
        """Rows of channel sums per image the fused squeeze writes; 0 = no such kernel. Asked of the library itself
        (csrc/dwconv.cu dwconv3x3_strips_c) so the plan and the kernel cannot disagree."""
        from . import lib as _lib
        return int(_lib.load().b200_dwconv3x3_squeeze_rows(int(c), int(ho), int(wo), int(stride)))

    def dwconv(self, x: T, weight, bias, stride, act, out: Optional[T] = None, name="", squeeze=False):
        """squeeze=True: also returns the fp32 strip-sum buffer [strips][cphys] feeding se_gate (fused SE squeeze)."""
        c = weight.shape[0]
        assert weight.shape[1] == 1 and weight.shape[2] == 3 and c == x.c
        ho, wo = (x.h - 1) // stride + 1, (x.w - 1) // stride + 1
        if out is None:
            ob = self.new_buf(ho, wo, x.cphys)
            out = T(ob, 0, c, ho, wo, x.cphys, x.cmap)      # same physical channel arrangement as the input
        assert out.cphys == x.cphys and np.array_equal(out.cmap, x.cmap)
        cp = x.cphys
        wp = torch.zeros(9, cp)
        wp[:, torch.as_tensor(x.cmap)] = weight.reshape(c, 9).t()
        bp = torch.zeros(cp)
        bp[torch.as_tensor(x.cmap)] = bias
        w_off = self.add_weights(wp.numpy().astype(np.float32))
        b_off = self.add_weights(bp.numpy().astype(np.float32))
        if squeeze:
            strips = self.dw_strips(cp, ho, wo, stride)
            assert strips > 0
            sums = self.new_buf(1, strips, cp, dtype=1)
            self._op(OP_DWCONV, [x.buf, x.coff, cp, out.buf, out.coff, cp, -1, 0, 3, stride, 1, act, sums], w_off, b_off,
                     name=name)
            return out, (sums, strips)
        self._op(OP_DWCONV, [x.buf, x.coff, cp, out.buf, out.coff, cp, -1, 0, 3, stride, 1, act], w_off, b_off, name=name)
        return out

    def se_gate(self, x: T, sums, w1, b1, w2, b2, name=""):
        """Fused SE gate from the depthwise strip sums of x: sigmoid(fc2(silu(fc1(mean)))) -> fp32 vector buffer [cphys].
        w1 fp32 [S, C], w2 fp32 [C, S] (torchvision's fc1 / fc2 1x1 conv weights, flattened)."""
        sums_buf, strips = sums
        sq, c = w1.shape
        assert c == x.c and tuple(w2.shape) == (c, sq)
        cp, sp = x.cphys, pad8(sq)
        cm = torch.as_tensor(x.cmap)
        w1p = torch.zeros(sp, cp)
        w1p[:sq, cm] = w1
        b1p = torch.zeros(sp)
        b1p[:sq] = b1
        w2t = torch.zeros(sp, cp)
        w2t[:sq, cm] = w2.t()
        b2p = torch.zeros(cp)
        b2p[cm] = b2
        out_buf = self.new_buf(1, 1, cp, dtype=1)
        offs = [self.add_weights(a.numpy().astype(np.float32)) for a in (w1p, b1p, w2t, b2p)]
        inv = struct.unpack("<i", struct.pack("<f", 1.0 / (x.h * x.w)))[0]
        self._op(OP_SE_GATE, [sums_buf, strips, cp, out_buf, 0, cp, -1, 0, sp, 0, 0, 0, inv], *offs, name=name)
        return T(out_buf, 0, x.c, 1, 1, cp, x.cmap)

    def maxpool(self, x: T, k, stride, pad, out: Optional[T] = None, name=""):
        ho, wo = (x.h + 2 * pad - k) // stride + 1, (x.w + 2 * pad - k) // stride + 1
        if out is None:
            out = T(self.new_buf(ho, wo, x.cphys), 0, x.c, ho, wo, x.cphys, x.cmap)
        assert out.cphys == x.cphys
        self._op(OP_MAXPOOL, [x.buf, x.coff, x.cphys, out.buf, out.coff, x.cphys, -1, 0, k, stride, pad, 0], name=name)
        return out

    def avgpool(self, x: T, name=""):
        """Global average pool -> fp32 vector buffer [c]."""
        out_buf = self.new_buf(1, 1, x.cphys, dtype=1)
        self._op(OP_AVGPOOL, [x.buf, x.coff, x.cphys, out_buf, 0, x.cphys], name=name)
        return T(out_buf, 0, x.c, 1, 1, x.cphys, x.cmap)

    def fc(self, x: T, weight, bias, act=ACT_NONE, name=""):
        """x: fp32 vector buffer; weight fp32 [J, K]."""
        j, k = weight.shape
        assert k == x.c and self.bufs[x.buf][3] == 1
        kp = self.bufs[x.buf][2]
        jp = pad8(j)
        wp = torch.zeros(jp, kp)
        wp[:j, torch.as_tensor(x.cmap)] = weight
        bp = torch.zeros(jp)
        bp[:j] = bias
        out_buf = self.new_buf(1, 1, jp, dtype=1)
        w_off = self.add_weights(wp.numpy().astype(np.float32))
        b_off = self.add_weights(bp.numpy().astype(np.float32))
        one = struct.unpack("<i", struct.pack("<f", 1.0))[0]
        self._op(OP_FC, [x.buf, 0, kp, out_buf, 0, jp, -1, 0, 0, 0, 0, act, one], w_off, b_off, name=name)
        return T(out_buf, 0, j, 1, 1, jp)

    def se_excite(self, x: T, sums, w1, b1, w2, b2, name=""):
        """SE block after the fused squeeze, in two ops: hidden = silu(fc1(mean)) [sp] fp32, then
        out = x * sigmoid(fc2(hidden)) with the gate computed per channel slice inside the scale kernel.
        w1 fp32 [S, C], w2 fp32 [C, S] (torchvision's fc1 / fc2 1x1 conv weights, flattened)."""
        sums_buf, rows = sums
        sq, c = w1.shape
        assert c == x.c and tuple(w2.shape) == (c, sq) and sq <= 256
        cp, sp = x.cphys, pad8(sq)
        cm = torch.as_tensor(x.cmap)
        w1p = torch.zeros(sp, cp)
        w1p[:sq, cm] = w1
        b1p = torch.zeros(sp)
        b1p[:sq] = b1
        w2t = torch.zeros(sp, cp)
        w2t[:sq, cm] = w2.t()
        b2p = torch.zeros(cp)
        b2p[cm] = b2
        hid_buf = self.new_buf(1, 1, sp, dtype=1)
        offs = [self.add_weights(a.numpy().astype(np.float32)) for a in (w1p, b1p, w2t, b2p)]
        inv = struct.unpack("<i", struct.pack("<f", 1.0 / (x.h * x.w)))[0]
        self._op(OP_SE_HIDDEN, [sums_buf, rows, cp, hid_buf, 0, sp, -1, 0, sp, 0, 0, 0, inv], offs[0], offs[1],
                 name=name + "hidden")
        out = T(self.new_buf(x.h, x.w, cp), 0, x.c, x.h, x.w, cp, x.cmap)
        self._op(OP_SE_SCALE, [x.buf, x.coff, cp, out.buf, out.coff, cp, -1, 0, sp, 0, 0, 0, hid_buf], offs[2], offs[3],
                 name=name + "scale")
        return out

    def scale(self, x: T, gate: T, name=""):
        assert gate.cphys == x.cphys and self.bufs[gate.buf][2] == x.cphys
        out = T(self.new_buf(x.h, x.w, x.cphys), 0, x.c, x.h, x.w, x.cphys, x.cmap)
        self._op(OP_SCALE, [x.buf, x.coff, x.cphys, out.buf, out.coff, x.cphys, -1, 0, 0, 0, 0, 0, gate.buf], name=name)
        return out

    def shuffle(self, a: T, b: T, name=""):
        """ShuffleNetV2 cat + channel_shuffle(groups=2): logical out[2i] = a[i], out[2i+1] = b[i]. The result is stored
        as two halves [0, cp) and [cp, 2cp) (cp = line_channels(c)) holding logical [0, c) and [c, 2c): chunk(2) of the next unit
        is then two 16-byte aligned slices."""
        assert a.c == b.c and (a.h, a.w) == (b.h, b.w)
        assert np.array_equal(a.cmap, np.arange(a.c)) and np.array_equal(b.cmap, np.arange(b.c))
        c, cp = a.c, line_channels(a.c)
        ob = self.new_buf(a.h, a.w, 2 * cp)
        cmap = np.concatenate([np.arange(c), cp + np.arange(c)])
        self._op(OP_SHUFFLE, [a.buf, a.coff, a.cphys, ob, 0, 2 * cp, -1, 0, 0, 0, 0, 0, b.buf, b.coff, c, cp], name=name)
        return T(ob, 0, 2 * c, a.h, a.w, 2 * cp, cmap)

    def conv_shuffle(self, x: T, weight, bias, other: T, name=""):
        """relu(conv1x1(x)) interleaved with `other` by the conv's own epilogue: the result is shuffle(other, conv) in the
        halves layout of `shuffle` (csrc/conv_gemm.cu, ConvSpec::interleave_c). The weights are laid out in cp "slots":
        output channel i < c/2 in slot i, i >= c/2 in slot cp/2 + (i - c/2); slot s lands on physical channel 2 s + 1 of
        the output and other's channel of the same slot on 2 s. Returns None when the shapes do not allow it."""
        c, cin = weight.shape[:2]
        cp = line_channels(c)
        if not (weight.shape[2:] == (1, 1) and other.c == c and other.cphys == cp and c % 2 == 0 and x.cphys % 8 == 0
                and (cp == 64 or cp % 128 == 0) and (x.h, x.w) == (other.h, other.w)
                and np.array_equal(other.cmap, np.arange(c))):
            return None
        assert cin == x.c
        slot = np.concatenate([np.arange(c // 2), cp // 2 + np.arange(c - c // 2)])
        wp = torch.zeros(cp, x.cphys)
        wfull = torch.zeros(c, x.cphys)
        wfull[:, torch.as_tensor(x.cmap)] = weight.reshape(c, cin)
        wp[torch.as_tensor(slot)] = wfull
        bp = torch.zeros(cp)
        bp[torch.as_tensor(slot)] = bias
        w_off = self.add_weights(self.bf16_bits(wp))
        b_off = self.add_weights(bp.numpy().astype(np.float32))
        ob = self.new_buf(x.h, x.w, 2 * cp)
        pr = [x.buf, x.coff, x.cphys, ob, 0, 2 * cp, other.buf, other.coff, 1, 1, 0, ACT_RELU, 0, 0, 0, 0, 0, 0, c]
        self._op(OP_CONV, pr, w_off, b_off, name=name)
        return T(ob, 0, 2 * c, x.h, x.w, 2 * cp, np.concatenate([np.arange(c), cp + np.arange(c)]))

    @staticmethod
    def chunk2(x: T):
        """x.chunk(2, dim=1) of a halves-layout tensor: two aligned slices, no copy."""
        c, cp = x.c // 2, x.cphys // 2
        assert np.array_equal(x.cmap, np.concatenate([np.arange(c), cp + np.arange(c)]))
        return T(x.buf, x.coff, c, x.h, x.w, cp), T(x.buf, x.coff + cp, c, x.h, x.w, cp)

    def copy(self, x: T, out: T, name=""):
        assert x.cphys == out.cphys
        self._op(OP_COPY, [x.buf, x.coff, x.cphys, out.buf, out.coff, x.cphys], name=name)
        return out

    def upsample2x(self, x: T, out: T, name=""):
        assert x.cphys == out.cphys
        self._op(OP_UPSAMPLE, [x.buf, x.coff, x.cphys, out.buf, out.coff, x.cphys], name=name)
        return out

    def yolo_decode(self, hb: T, hc: T, wb, bb, wc, bc, out_buf, a_off, a_total, stride, name=""):
        """Final 1x1 convs of one Detect level (fp32) + DFL + dist2bbox + sigmoid into the [84][A] output buffer."""
        assert hb.c == 64 and hc.c == 80 and wb.shape == (64, 64) and wc.shape == (80, 80)
        w_off = self.add_weights(wb.numpy().astype(np.float32))
        b_off = self.add_weights(bb.numpy().astype(np.float32))
        w2_off = self.add_weights(wc.numpy().astype(np.float32))
        b2_off = self.add_weights(bc.numpy().astype(np.float32))
        p = [hb.buf, hb.coff, 64, out_buf, 0, 0, -1, 0, 0, 0, 0, 0, hc.buf, hc.coff, 80, a_off, int(stride), a_total]
        self._op(OP_YOLO_DECODE, p, w_off, b_off, w2_off, b2_off, name=name)

    def slice(self, buf, coff, c, h, w):
        return T(buf, coff, c, h, w)

    def set_output(self, t: T):
        assert self.bufs[t.buf][3] == 1, "model output must be an fp32 vector buffer"
        assert np.array_equal(t.cmap, np.arange(t.c))
        self.out_buf = t.buf
        self.out_dim = self.bufs[t.buf][2]
        self.out_logical = t.c

    # ---- serialisation ----
    def to_blob(self) -> bytes:
        assert self.out_buf >= 0
        tables = HEADER_SIZE + len(self.bufs) * struct.calcsize(BUF_FMT) + len(self.ops) * 128
        w_offset = (tables + 255) // 256 * 256
        hdr = struct.pack(HEADER_FMT, b"B200PLAN", 1, len(self.bufs), len(self.ops), 0, self.kind.encode()[:31],
                          self.in_h, self.in_w, self.in_buf, self.out_buf, self.out_dim, self.pre_fill, *self.mean,
                          *self.std, w_offset, len(self.weights))
        out = bytearray(hdr)
        for b in self.bufs:
            out += struct.pack(BUF_FMT, *b)
        for typ, p, w, b, w2, b2 in self.ops:
            out += struct.pack(OP_FMT, typ, *p, w, b, w2, b2)
        out += b"\0" * (w_offset - len(out))
        out += self.weights
        return bytes(out)


def parse_blob(blob: bytes):
    """Inverse of PlanBuilder.to_blob: returns (header dict, bufs, ops, weights bytes)."""
    f = struct.unpack_from(HEADER_FMT, blob, 0)
    hdr = dict(magic=f[0], version=f[1], n_bufs=f[2], n_ops=f[3], kind=f[5].split(b"\0")[0].decode(), in_h=f[6],
               in_w=f[7], in_buf=f[8], out_buf=f[9], out_dim=f[10], pre_fill=f[11], mean=f[12:15], std=f[15:18],
               weights_offset=f[18], weights_bytes=f[19])
    off = HEADER_SIZE
    bufs = []
    for _ in range(hdr["n_bufs"]):
        bufs.append(struct.unpack_from(BUF_FMT, blob, off))
        off += struct.calcsize(BUF_FMT)
    ops = []
    for _ in range(hdr["n_ops"]):
        r = struct.unpack_from(OP_FMT, blob, off)
        ops.append(dict(type=r[0], p=list(r[1:24]), w_off=r[24], b_off=r[25], w2_off=r[26], b2_off=r[27]))
        off += 128
    weights = blob[hdr["weights_offset"]:hdr["weights_offset"] + hdr["weights_bytes"]]
    return hdr, bufs, ops, weights


# ================================================================================================ model lowerings

def lower_resnet(model, in_hw=(224, 224)) -> PlanBuilder:
    """torchvision ResNet (Bottleneck, stride on the 3x3 = v1.5): tv models/resnet.py:108-165, 266-285."""
    pb = PlanBuilder("resnet50", in_hw[0], in_hw[1])
    w, b = fold_bn(model.conv1, model.bn1)
    hs, ws = (in_hw[0] - 1) // 2 + 1, (in_hw[1] - 1) // 2 + 1
    if pb.fuse_stem_pool and ws <= 128 and hs % 2 == 0 and ws % 2 == 0 and in_hw[1] % 16 == 0:
        x = pb.stem_pool(pb.input(), w, b, ACT_RELU, name="conv1+maxpool")
    else:
        x = pb.stem(pb.input(), w, b, 2, 3, ACT_RELU, name="conv1")
        x = pb.maxpool(x, 3, 2, 1, name="maxpool")
    for li, layer in enumerate([model.layer1, model.layer2, model.layer3, model.layer4], 1):
        for bi, blk in enumerate(layer):
            nm = "layer%d.%d." % (li, bi)
            identity = x
            w, b = fold_bn(blk.conv1, blk.bn1)
            o = pb.conv(x, w, b, 1, 0, ACT_RELU, name=nm + "conv1")
            w, b = fold_bn(blk.conv2, blk.bn2)
            o = pb.conv(o, w, b, blk.conv2.stride[0], 1, ACT_RELU, name=nm + "conv2")
            w, b = fold_bn(blk.conv3, blk.bn3)
            if blk.downsample is not None and pb.fuse_downsample and o.cphys % 64 == 0:
                # conv3 and the downsample branch are both 1x1 convs that are summed: one GEMM over [conv2 out | block in]
                wd, bd = fold_bn(blk.downsample[0], blk.downsample[1])
                x = pb.conv_dual(o, w, b, x, wd, bd, blk.downsample[0].stride[0], ACT_RELU, name=nm + "conv3+downsample")
                continue
            if blk.downsample is not None:
                wd, bd = fold_bn(blk.downsample[0], blk.downsample[1])
                identity = pb.conv(x, wd, bd, blk.downsample[0].stride[0], 0, ACT_NONE, name=nm + "downsample")
            x = pb.conv(o, w, b, 1, 0, ACT_RELU, res=identity, name=nm + "conv3")
    v = pb.avgpool(x, name="avgpool")
    y = pb.fc(v, model.fc.weight.detach().float(), model.fc.bias.detach().float(), name="fc")
    pb.set_output(y)
    return pb


def lower_shufflenet(model, in_hw=(224, 224)) -> PlanBuilder:
    """torchvision ShuffleNetV2 x1.0: tv models/shufflenetv2.py:43-104 (InvertedResidual), 149-163 (_forward_impl)."""
    pb = PlanBuilder("shufflenet_v2_x1_0", in_hw[0], in_hw[1])
    w, b = fold_bn(model.conv1[0], model.conv1[1])
    hs, ws = (in_hw[0] - 1) // 2 + 1, (in_hw[1] - 1) // 2 + 1
    if pb.fuse_stem_pool and pb.toeplitz_stem and ws <= 128 and hs % 2 == 0 and ws % 2 == 0 and in_hw[1] % 16 == 0:
        x = pb.stem_pool(pb.input(), w, b, ACT_RELU, name="conv1+maxpool")
    else:
        x = pb.stem(pb.input(), w, b, 2, 1, ACT_RELU, name="conv1")
        x = pb.maxpool(x, 3, 2, 1, name="maxpool")
    for sname in ("stage2", "stage3", "stage4"):
        for ui, unit in enumerate(getattr(model, sname)):
            nm = "%s.%d." % (sname, ui)
            b2 = unit.branch2
            if unit.stride == 1:
                x1, x2 = pb.chunk2(x)
                a = x1
                src = x2
            else:
                b1 = unit.branch1
                w, b = fold_bn(b1[0], b1[1])
                t = pb.dwconv(x, w, b, unit.stride, ACT_NONE, name=nm + "branch1.dw")
                w, b = fold_bn(b1[2], b1[3])
                a = pb.conv(t, w, b, 1, 0, ACT_RELU, name=nm + "branch1.pw")
                src = x
            w, b = fold_bn(b2[0], b2[1])
            t = pb.conv(src, w, b, 1, 0, ACT_RELU, name=nm + "branch2.pw1")
            w, b = fold_bn(b2[3], b2[4])
            t = pb.dwconv(t, w, b, unit.stride, ACT_NONE, name=nm + "branch2.dw")
            w, b = fold_bn(b2[5], b2[6])
            fused = pb.conv_shuffle(t, w, b, a, name=nm + "branch2.pw2+shuffle") if pb.fuse_shuffle else None
            if fused is not None:
                x = fused
                continue
            t = pb.conv(t, w, b, 1, 0, ACT_RELU, name=nm + "branch2.pw2")
            x = pb.shuffle(a, t, name=nm + "shuffle")
    w, b = fold_bn(model.conv5[0], model.conv5[1])
    x = pb.conv(x, w, b, 1, 0, ACT_RELU, name="conv5")
    v = pb.avgpool(x, name="mean")
    y = pb.fc(v, model.fc.weight.detach().float(), model.fc.bias.detach().float(), name="fc")
    pb.set_output(y)
    return pb


def lower_efficientnet_v2(model, in_hw=(224, 224), fuse_se=None) -> PlanBuilder:
    """torchvision EfficientNetV2-S: tv models/efficientnet.py:105-170 (MBConv), 172-231 (FusedMBConv), 333-345;
    SqueezeExcitation: tv ops/misc.py. stochastic_depth is the identity in eval mode."""
    from torchvision.models.efficientnet import FusedMBConv, MBConv
    pb = PlanBuilder("efficientnet_v2_s", in_hw[0], in_hw[1])
    if fuse_se is None:      # depthwise conv writes the squeeze sums; one gate kernel replaces avgpool + fc1 + fc2
        fuse_se = os.environ.get("B200_FUSE_SE", "1") != "0"
    # hidden + (fc2 + scale) instead of gate + scale: measured equal within noise on B200 (3.87 vs 3.81 ms @128), off
    split_se = os.environ.get("B200_SPLIT_SE", "0") == "1"
    # x * gate inside the project conv instead of a scale pass: bit-exact but slower on B200 (the in-smem rescale competes
    # with the tensor core for shared-memory bandwidth: 73 vs 22 + 18 us per block), so off by default
    fold_scale = os.environ.get("B200_FOLD_SE_SCALE", "0") == "1"
    feats = model.features
    w, b = fold_bn(feats[0][0], feats[0][1])
    x = pb.stem(pb.input(), w, b, 2, 1, ACT_SILU, name="features.0")
    for si in range(1, len(feats) - 1):
        for bi, blk in enumerate(feats[si]):
            nm = "features.%d.%d." % (si, bi)
            res = x if blk.use_res_connect else None
            if isinstance(blk, FusedMBConv):
                layers = list(blk.block)
                if len(layers) == 1:      # expand_ratio == 1: conv3x3 + BN + SiLU, then (+ input)
                    cna = layers[0]
                    w, b = fold_bn(cna[0], cna[1])
                    x = pb.conv(x, w, b, cna[0].stride[0], 1, ACT_SILU, res=res, res_after_act=True, name=nm + "fused")
                else:
                    cna, proj = layers
                    w, b = fold_bn(cna[0], cna[1])
                    t = pb.conv(x, w, b, cna[0].stride[0], 1, ACT_SILU, name=nm + "expand3x3")
                    w, b = fold_bn(proj[0], proj[1])
                    x = pb.conv(t, w, b, 1, 0, ACT_NONE, res=res, name=nm + "project")
            elif isinstance(blk, MBConv):
                layers = list(blk.block)
                exp, dw, se, proj = layers
                w, b = fold_bn(exp[0], exp[1])
                t = pb.conv(x, w, b, 1, 0, ACT_SILU, name=nm + "expand")
                w, b = fold_bn(dw[0], dw[1])
                st = dw[0].stride[0]
                if fuse_se and pb.dw_strips(t.cphys, (t.h - 1) // st + 1, (t.w - 1) // st + 1, st) > 0:
                    t, sums = pb.dwconv(t, w, b, st, ACT_SILU, name=nm + "dw", squeeze=True)
                    fcs = (se.fc1.weight.detach().float().flatten(1), se.fc1.bias.detach().float(),
                           se.fc2.weight.detach().float().flatten(1), se.fc2.bias.detach().float())
                    if split_se:
                        t = pb.se_excite(t, sums, *fcs, name=nm + "se.")
                        s = None
                    else:
                        s = pb.se_gate(t, sums, *fcs, name=nm + "se.gate")
                else:
                    t = pb.dwconv(t, w, b, st, ACT_SILU, name=nm + "dw")
                    s = pb.avgpool(t, name=nm + "se.squeeze")
                    s = pb.fc(s, se.fc1.weight.detach().float().flatten(1), se.fc1.bias.detach().float(), ACT_SILU,
                              name=nm + "se.fc1")
                    s = pb.fc(s, se.fc2.weight.detach().float().flatten(1), se.fc2.bias.detach().float(), ACT_SIGMOID,
                              name=nm + "se.fc2")
                w, b = fold_bn(proj[0], proj[1])
                if s is not None and fold_scale:
                    # SE excitation folded into the project conv: its A tile is scaled in shared memory, no scale pass
                    x = pb.conv(t, w, b, 1, 0, ACT_NONE, res=res, name=nm + "se.scale+project", in_scale=s)
                    continue
                if s is not None:
                    t = pb.scale(t, s, name=nm + "se.scale")
                x = pb.conv(t, w, b, 1, 0, ACT_NONE, res=res, name=nm + "project")
            else:
                raise NotImplementedError(type(blk))
    last = feats[-1]
    w, b = fold_bn(last[0], last[1])
    x = pb.conv(x, w, b, 1, 0, ACT_SILU, name="features.last")
    v = pb.avgpool(x, name="avgpool")
    fc = model.classifier[1]
    y = pb.fc(v, fc.weight.detach().float(), fc.bias.detach().float(), name="classifier")
    pb.set_output(y)
    return pb


def lower_yolov8n(model, in_hw=(640, 640)) -> PlanBuilder:
    """YOLOv8n (ultralytics yolov8.yaml, scale n; see yolov8n_arch.py). Accepts either our plain-PyTorch module (layers
    m0..m22) or an ultralytics DetectionModel (`.model[i]`). Concats are written in place: every producer of a concat
    operand stores straight into its channel slice of the concat buffer."""
    def L(i):
        return model.model[i] if hasattr(model, "model") else getattr(model, "m%d" % i)

    pb = PlanBuilder("yolov8n", in_hw[0], in_hw[1], mean=(0.0, 0.0, 0.0), std=(1.0, 1.0, 1.0), pre_fill=114)

    def conv_bn_act(x, m, out=None, res=None, name=""):
        w, b = fold_bn(m.conv, m.bn)
        k, s = m.conv.kernel_size[0], m.conv.stride[0]
        return pb.conv(x, w, b, s, k // 2, ACT_SILU, res=res, out=out, res_after_act=res is not None, name=name)

    def c2f(x, m, out=None, name=""):
        c, n = m.c, len(m.m)
        cat = pb.new_buf(x.h, x.w, (2 + n) * c)
        conv_bn_act(x, m.cv1, out=T(cat, 0, 2 * c, x.h, x.w), name=name + ".cv1")
        last = T(cat, c, c, x.h, x.w)
        for i, bt in enumerate(m.m):
            t = conv_bn_act(last, bt.cv1, name="%s.m.%d.cv1" % (name, i))
            dst = T(cat, (2 + i) * c, c, x.h, x.w)
            conv_bn_act(t, bt.cv2, out=dst, res=last if bt.add else None, name="%s.m.%d.cv2" % (name, i))
            last = dst
        return conv_bn_act(T(cat, 0, (2 + n) * c, x.h, x.w), m.cv2, out=out, name=name + ".cv2")

    H, W = in_hw
    w, b = fold_bn(L(0).conv, L(0).bn)
    x = pb.stem(pb.input(), w, b, 2, 1, ACT_SILU, name="0")                    # 320x320x16
    x = conv_bn_act(x, L(1), name="1")                                         # 160x160x32
    x = c2f(x, L(2), name="2")
    x = conv_bn_act(x, L(3), name="3")                                         # 80x80x64
    cat14 = pb.new_buf(H // 8, W // 8, 128 + 64)                                # [up(12) | p3]
    p3 = c2f(x, L(4), out=T(cat14, 128, 64, H // 8, W // 8), name="4")
    x = conv_bn_act(p3, L(5), name="5")                                        # 40x40x128
    cat11 = pb.new_buf(H // 16, W // 16, 256 + 128)                             # [up(9) | p4]
    p4 = c2f(x, L(6), out=T(cat11, 256, 128, H // 16, W // 16), name="6")
    x = conv_bn_act(p4, L(7), name="7")                                        # 20x20x256
    x = c2f(x, L(8), name="8")
    # SPPF
    sp = L(9)
    c_ = sp.cv1.conv.out_channels
    spcat = pb.new_buf(H // 32, W // 32, 4 * c_)
    y = conv_bn_act(x, sp.cv1, out=T(spcat, 0, c_, H // 32, W // 32), name="9.cv1")
    for i in range(3):
        y = pb.maxpool(y, 5, 1, 2, out=T(spcat, (i + 1) * c_, c_, H // 32, W // 32), name="9.m%d" % i)
    cat20 = pb.new_buf(H // 32, W // 32, 128 + 256)                             # [conv19 | p5]
    p5 = conv_bn_act(T(spcat, 0, 4 * c_, H // 32, W // 32), sp.cv2, out=T(cat20, 128, 256, H // 32, W // 32), name="9.cv2")
    # head
    pb.upsample2x(p5, T(cat11, 0, 256, H // 16, W // 16), name="10")
    cat17 = pb.new_buf(H // 16, W // 16, 64 + 128)                              # [conv16 | h12]
    h12 = c2f(T(cat11, 0, 384, H // 16, W // 16), L(12), out=T(cat17, 64, 128, H // 16, W // 16), name="12")
    pb.upsample2x(h12, T(cat14, 0, 128, H // 8, W // 8), name="13")
    h15 = c2f(T(cat14, 0, 192, H // 8, W // 8), L(15), name="15")              # P3 out, 64 ch
    conv_bn_act(h15, L(16), out=T(cat17, 0, 64, H // 16, W // 16), name="16")
    h18 = c2f(T(cat17, 0, 192, H // 16, W // 16), L(18), name="18")            # P4 out, 128 ch
    conv_bn_act(h18, L(19), out=T(cat20, 0, 128, H // 32, W // 32), name="19")
    h21 = c2f(T(cat20, 0, 384, H // 32, W // 32), L(21), name="21")            # P5 out, 256 ch
    # Detect
    det = L(22)
    feats = [h15, h18, h21]
    a_total = sum(f.h * f.w for f in feats)
    out_buf = pb.new_buf(1, 1, 84 * a_total, dtype=1)
    a_off = 0
    for i, f in enumerate(feats):
        hb = conv_bn_act(conv_bn_act(f, det.cv2[i][0], name="22.cv2.%d.0" % i), det.cv2[i][1], name="22.cv2.%d.1" % i)
        hc = conv_bn_act(conv_bn_act(f, det.cv3[i][0], name="22.cv3.%d.0" % i), det.cv3[i][1], name="22.cv3.%d.1" % i)
        fb, fc_ = det.cv2[i][2], det.cv3[i][2]
        pb.yolo_decode(hb, hc, fb.weight.detach().float().flatten(1), fb.bias.detach().float(),
                       fc_.weight.detach().float().flatten(1), fc_.bias.detach().float(), out_buf, a_off, a_total,
                       int(det.stride[i]), name="22.decode.%d" % i)
        a_off += f.h * f.w
    pb.set_output(T(out_buf, 0, 84 * a_total, 1, 1, 84 * a_total))
    return pb


LOWERINGS = {"yolov8n": lower_yolov8n, "resnet50": lower_resnet, "shufflenet_v2_x1_0": lower_shufflenet, "efficientnet_v2_s": lower_efficientnet_v2}


def lower(kind: str, model, **kw) -> PlanBuilder:
    if kind not in LOWERINGS:
        raise KeyError("no lowering for model kind %r (have %s)" % (kind, sorted(LOWERINGS)))
    model.eval()
    with torch.no_grad():
        return LOWERINGS[kind](model, **kw)
