"""ORACLE (test infrastructure only — never imported by the product path).

CPU interpreter of a libb200infer plan blob in plain PyTorch fp32.  It replays the fused op list with the same storage
precision as the CUDA engine (bf16 activations / conv weights, fp32 accumulation, fp32 heads), so it serves two checks:
  * lowering check (no GPU needed): interpreter(plan(model)) ~= model(x) of the torchvision fp32 reference
    (scheduler.py:449-450 `outputs = model(inputs)`) within the north-star tolerance;
  * kernel check (GPU): every activation buffer of the CUDA engine ~= the interpreter's buffer of the same id.
Parity of this interpreter is pinned by tests/test_lowering.py against torchvision's own forward.
"""
import struct

import numpy as np
import torch
import torch.nn.functional as F

OP_CONV, OP_STEM, OP_MAXPOOL, OP_AVGPOOL, OP_FC, OP_DWCONV, OP_SCALE, OP_COPY, OP_UPSAMPLE, OP_SHUFFLE, \
    OP_YOLO_DECODE, OP_GATHER_CONV, OP_SE_GATE, OP_SE_HIDDEN, OP_SE_SCALE = range(1, 16)


def _bf16(x):
    return x.to(torch.bfloat16).float()


def _act(y, a):
    if a == 1:
        return F.relu(y)
    if a == 2:
        return F.silu(y)
    if a == 3:
        return torch.sigmoid(y)
    return y


def stem_cout_tile(cout):
    return 32 if cout % 32 == 0 else (24 if cout % 24 == 0 else 16)


class PlanInterpreter:
    def __init__(self, blob, parse_blob):
        self.hdr, self.bufs, self.ops, self.weights = parse_blob(blob)
        self.wbytes = np.frombuffer(self.weights, dtype=np.uint8)

    def _f32(self, off, n):
        return torch.from_numpy(self.wbytes[off:off + 4 * n].view(np.float32).copy())

    def _bf16w(self, off, n):
        bits = self.wbytes[off:off + 2 * n].view(np.uint16).astype(np.int32) << 16
        return torch.from_numpy(bits.view(np.float32).copy())

    def run(self, x_nhwc4, keep_all=False):
        """x_nhwc4: float tensor [N, H, W, 4] holding bf16-representable values. Returns (out [N, out_dim], buffers)."""
        n = x_nhwc4.shape[0]
        B = [None] * len(self.bufs)
        for i, (h, w, c, dt) in enumerate(self.bufs):
            B[i] = torch.zeros(n, h, w, c)
        B[self.hdr["in_buf"]] = _bf16(x_nhwc4.float())
        for op in self.ops:
            t, p = op["type"], op["p"]
            if t == OP_CONV:
                cin, cout, k, s, pad, act = p[2], p[5], p[8], p[9], p[10], p[11]
                x = B[p[0]][..., p[1]:p[1] + cin]
                if p[17] > 0:      # SE gate folded into the conv input: x * gate, rounded to bf16 like the scale op's output
                    x = _bf16(x * B[p[17] - 1].view(n, 1, 1, -1)[..., :cin])
                x = x.permute(0, 3, 1, 2)
                if p[18] > 0:      # ShuffleNetV2 unit tail: channel_shuffle(cat(other, relu(conv)), 2) in halves layout;
                    c, cp = p[18], cout // 2          # weights in cp slots (channel i < c/2 in slot i, else cp/2 + i - c/2)
                    w = self._bf16w(op["w_off"], cp * cin).view(cp, cin)
                    y = _bf16(_act(F.conv2d(x, w[:, :, None, None], self._f32(op["b_off"], cp)), act)).permute(0, 2, 3, 1)
                    other = B[p[6]][..., p[7]:p[7] + cp]
                    xs = torch.zeros_like(y)
                    xs[..., :c // 2] = other[..., :c // 2]
                    xs[..., cp // 2:cp // 2 + c // 2] = other[..., c // 2:c]
                    B[p[3]][..., p[4]:p[4] + 2 * cp] = torch.stack([xs, y], dim=-1).reshape(*y.shape[:-1], 2 * cp)
                    continue
                b = self._f32(op["b_off"], cout)
                rows = p[20] if p[20] > 0 else cout      # weights stored as [rows][k*k][kpad], zero padded (0 = exact)
                if p[13] > 0:      # K-concatenated 1x1 conv: W[:, :cin] . x + W[:, cin:] . x2[::s2, ::s2]
                    cin2, s2 = p[15], p[16]
                    kpad = p[19] if p[19] > 0 else cin + cin2
                    wk = self._bf16w(op["w_off"], rows * kpad).view(rows, kpad)[:cout]
                    x2 = B[p[13] - 1][..., p[14]:p[14] + cin2].permute(0, 3, 1, 2)
                    y = F.conv2d(x, wk[:, :cin, None, None], b) + \
                        F.conv2d(x2, wk[:, cin:cin + cin2, None, None], None, stride=s2)
                else:
                    kpad = p[19] if p[19] > 0 else cin
                    w = self._bf16w(op["w_off"], rows * k * k * kpad).view(rows, k, k, kpad)[:cout, :, :, :cin]
                    y = F.conv2d(x, w.permute(0, 3, 1, 2), b, stride=s, padding=pad)
                if p[6] >= 0:
                    r = B[p[6]][..., p[7]:p[7] + cout].permute(0, 3, 1, 2)
                    y = _act(y, act) + r if p[12] else _act(y + r, act)
                else:
                    y = _act(y, act)
                B[p[3]][..., p[4]:p[4] + cout] = _bf16(y).permute(0, 2, 3, 1)
            elif t == OP_GATHER_CONV:
                cin, cout, k, s, pad, act = p[2], p[5], p[8], p[9], p[10], p[11]
                npad = 32 if cout <= 32 else (64 if cout <= 64 else (128 if cout <= 128 else 256))
                kpad = (k * k * (cin // 8) + 7) // 8 * 64
                x = B[p[0]][..., p[1]:p[1] + cin].permute(0, 3, 1, 2)
                wt = self._bf16w(op["w_off"], npad * kpad).view(npad, kpad)[:cout, :k * k * cin]
                w = wt.reshape(cout, k, k, cin).permute(0, 3, 1, 2)
                b = self._f32(op["b_off"], npad)[:cout]
                y = F.conv2d(x, w, b, stride=s, padding=pad)
                if p[6] >= 0:
                    r = B[p[6]][..., p[7]:p[7] + cout].permute(0, 3, 1, 2)
                    y = _act(y, act) + r if p[12] else _act(y + r, act)
                else:
                    y = _act(y, act)
                B[p[3]][..., p[4]:p[4] + cout] = _bf16(y).permute(0, 2, 3, 1)
            elif t == OP_STEM:
                cout, k, s, pad, act = p[5], p[8], p[9], p[10], p[11]
                ct = stem_cout_tile(cout)
                cpad = (cout + ct - 1) // ct * ct
                x = B[p[0]][..., :3].permute(0, 3, 1, 2)
                if p[12] in (2, 3):      # Toeplitz stem (3: fused with MaxPool2d(3,2,1)): bf16 image [k][kc][nout][8], element e of chunk c = pixel 2c + e//4, channel e%4
                    kc = 4 if k == 7 else 2
                    nout = 32 if cout <= 32 else 64
                    img = self._bf16w(op["w_off"], k * kc * nout * 8).view(k, kc, nout, 8)
                    w = torch.zeros(cout, 3, k, k)
                    for r in range(k):
                        for q in range(k):
                            wp_ = q + 1
                            w[:, :, r, q] = img[r, wp_ // 2, :cout, (wp_ % 2) * 4:(wp_ % 2) * 4 + 3]
                    b = self._f32(op["b_off"], nout)[:cout]
                elif p[12] == 1:      # tensor-core stem layout: bf16 [nout][kpad], K = r*row + (s+1)*4 + c
                    row, kpad = (32, 256) if k == 7 else (16, 64)
                    nout = 16 if cout <= 16 else (32 if cout <= 32 else 64)
                    wt = self._bf16w(op["w_off"], nout * kpad).view(nout, kpad)
                    w = torch.zeros(cout, 3, k, k)
                    for r in range(k):
                        for q in range(k):
                            w[:, :, r, q] = wt[:cout, r * row + (q + 1) * 4:r * row + (q + 1) * 4 + 3]
                    b = self._f32(op["b_off"], nout)[:cout]
                else:
                    w = self._f32(op["w_off"], k * k * 3 * cpad).view(k, k, 3, cpad)[..., :cout].permute(3, 2, 0, 1)
                    b = self._f32(op["b_off"], cpad)[:cout]
                y = _bf16(_act(F.conv2d(x, w, b, stride=s, padding=pad), act))
                if p[12] == 3:
                    y = F.max_pool2d(y, 3, 2, 1)
                B[p[3]][..., p[4]:p[4] + cout] = y.permute(0, 2, 3, 1)
            elif t == OP_DWCONV:
                c, s, act = p[2], p[9], p[11]
                x = B[p[0]][..., p[1]:p[1] + c].permute(0, 3, 1, 2)
                w = self._f32(op["w_off"], 9 * c).view(3, 3, c).permute(2, 0, 1).unsqueeze(1)
                b = self._f32(op["b_off"], c)
                y = F.conv2d(x, w, b, stride=s, padding=1, groups=c)
                y = _bf16(_act(y, act)).permute(0, 2, 3, 1)
                B[p[3]][..., p[4]:p[4] + c] = y
                if p[12] > 0:      # fused squeeze: per-strip channel sums of the stored outputs; the gate only uses their
                    B[p[12]].zero_()   # total, so the restatement puts the whole image in strip 0
                    B[p[12]].view(n, -1)[:, :c] = y.sum(dim=(1, 2))
            elif t == OP_SE_GATE:
                strips, c, sq = p[1], p[2], p[8]
                inv_hw = struct.unpack("<f", struct.pack("<i", p[12]))[0]
                mean = B[p[0]].view(n, -1)[:, :strips * c].view(n, strips, c).sum(dim=1) * inv_hw
                w1 = self._f32(op["w_off"], sq * c).view(sq, c)
                b1 = self._f32(op["b_off"], sq)
                w2t = self._f32(op["w2_off"], sq * c).view(sq, c)
                b2 = self._f32(op["b2_off"], c)
                hid = F.silu(mean @ w1.t() + b1)
                B[p[3]] = torch.sigmoid(hid @ w2t + b2).view(n, 1, 1, c)
            elif t == OP_SE_HIDDEN:
                rows, c, sq = p[1], p[2], p[8]
                inv_hw = struct.unpack("<f", struct.pack("<i", p[12]))[0]
                mean = B[p[0]].view(n, -1)[:, :rows * c].view(n, rows, c).sum(dim=1) * inv_hw
                w1 = self._f32(op["w_off"], sq * c).view(sq, c)
                b1 = self._f32(op["b_off"], sq)
                B[p[3]] = F.silu(mean @ w1.t() + b1).view(n, 1, 1, sq)
            elif t == OP_SE_SCALE:
                c, sq = p[2], p[8]
                w2t = self._f32(op["w_off"], sq * c).view(sq, c)
                b2 = self._f32(op["b_off"], c)
                g = torch.sigmoid(B[p[12]].view(n, -1)[:, :sq] @ w2t + b2).view(n, 1, 1, c)
                B[p[3]][..., p[4]:p[4] + c] = _bf16(B[p[0]][..., p[1]:p[1] + c] * g)
            elif t == OP_MAXPOOL:
                c, k, s, pad = p[2], p[8], p[9], p[10]
                x = B[p[0]][..., p[1]:p[1] + c].permute(0, 3, 1, 2)
                B[p[3]][..., p[4]:p[4] + c] = F.max_pool2d(x, k, s, pad).permute(0, 2, 3, 1)
            elif t == OP_AVGPOOL:
                c = p[2]
                x = B[p[0]][..., p[1]:p[1] + c]
                B[p[3]] = x.mean(dim=(1, 2)).view(n, 1, 1, c)
            elif t == OP_FC:
                k, j, act = p[2], p[5], p[11]
                scale = struct.unpack("<f", struct.pack("<i", p[12]))[0]
                x = B[p[0]].view(n, -1)[:, :k]
                w = self._f32(op["w_off"], j * k).view(j, k)
                b = self._f32(op["b_off"], j)
                B[p[3]] = _act((x @ w.t()) * scale + b, act).view(n, 1, 1, j)
            elif t == OP_SCALE:
                c = p[2]
                g = B[p[12]].view(n, 1, 1, -1)[..., :c]
                B[p[3]][..., p[4]:p[4] + c] = _bf16(B[p[0]][..., p[1]:p[1] + c] * g)
            elif t == OP_COPY:
                c = p[2]
                B[p[3]][..., p[4]:p[4] + c] = B[p[0]][..., p[1]:p[1] + c]
            elif t == OP_UPSAMPLE:
                c = p[2]
                x = B[p[0]][..., p[1]:p[1] + c]
                B[p[3]][..., p[4]:p[4] + c] = x.repeat_interleave(2, dim=1).repeat_interleave(2, dim=2)
            elif t == OP_SHUFFLE:
                c, cp = p[14], p[15]
                a = B[p[0]][..., p[1]:p[1] + c]
                b = B[p[12]][..., p[13]:p[13] + c]
                sh = torch.stack([a, b], dim=-1).reshape(*a.shape[:-1], 2 * c)   # channel_shuffle(cat(a, b), 2)
                o = B[p[3]]
                o[..., p[4]:p[4] + 2 * cp] = 0
                o[..., p[4]:p[4] + c] = sh[..., :c]
                o[..., p[4] + cp:p[4] + cp + c] = sh[..., c:]
            elif t == OP_YOLO_DECODE:
                hb = B[p[0]][..., p[1]:p[1] + 64]
                hc = B[p[12]][..., p[13]:p[13] + 80]
                a_off, stride, a_total = p[15], float(p[16]), p[17]
                wb = self._f32(op["w_off"], 64 * 64).view(64, 64)
                bb = self._f32(op["b_off"], 64)
                wc = self._f32(op["w2_off"], 80 * 80).view(80, 80)
                bc = self._f32(op["b2_off"], 80)
                h, w = hb.shape[1], hb.shape[2]
                box = (hb.reshape(n, h * w, 64) @ wb.t() + bb).view(n, h * w, 4, 16).softmax(3)
                dist = (box * torch.arange(16, dtype=torch.float32)).sum(3)            # [n, hw, 4] l t r b
                yy, xx = torch.meshgrid(torch.arange(h, dtype=torch.float32) + 0.5,
                                        torch.arange(w, dtype=torch.float32) + 0.5, indexing="ij")
                ax, ay = xx.reshape(-1), yy.reshape(-1)
                x1, y1, x2, y2 = ax - dist[..., 0], ay - dist[..., 1], ax + dist[..., 2], ay + dist[..., 3]
                cls = torch.sigmoid(hc.reshape(n, h * w, 80) @ wc.t() + bc)
                o = B[p[3]].view(n, 84, a_total)
                o[:, 0, a_off:a_off + h * w] = (x1 + x2) * 0.5 * stride
                o[:, 1, a_off:a_off + h * w] = (y1 + y2) * 0.5 * stride
                o[:, 2, a_off:a_off + h * w] = (x2 - x1) * stride
                o[:, 3, a_off:a_off + h * w] = (y2 - y1) * stride
                o[:, 4:, a_off:a_off + h * w] = cls.permute(0, 2, 1)
            else:
                raise NotImplementedError("op type %d" % t)
        out = B[self.hdr["out_buf"]].view(n, -1)
        return out, (B if keep_all else None)


def to_nhwc4(x_nchw):
    """[N,3,H,W] float -> [N,H,W,4] float with bf16-rounded values (what the collate kernel produces)."""
    n, c, h, w = x_nchw.shape
    out = torch.zeros(n, h, w, 4)
    out[..., :3] = _bf16(x_nchw.float()).permute(0, 2, 3, 1)
    return out
