"""YOLOv8n module definition in plain PyTorch (the container of weights that `lowering.lower("yolov8n", module)` consumes when
the `ultralytics` package is absent, and — run on the CPU in fp32 — the reference arm of the parity tests).

The reference repository names this model only through its image workload (`sample-models/yolov8/workload/`); the model
blobs are missing and nothing there runs it.  The architecture below restates ultralytics 8.0/8.1 `yolov8.yaml` at scale n
(depth 0.33, width 0.25): Conv = Conv2d + BN(eps 1e-3) + SiLU; C2f / Bottleneck / SPPF; Detect with reg_max 16, DFL
expectation, dist2bbox around (grid + 0.5), x stride, class sigmoid; output [B, 84, 8400].  Attribute names follow
ultralytics (conv/bn/act, cv1/cv2/m, cv2/cv3) so that a real ultralytics `DetectionModel.model` lowers the same way.
B200Model never calls these forward() methods.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


class Conv(nn.Module):
    """Conv2d(bias=False) + BatchNorm2d(eps=1e-3, momentum=0.03) + SiLU."""

    def __init__(self, c1, c2, k=1, s=1):
        super().__init__()
        self.conv = nn.Conv2d(c1, c2, k, s, k // 2, bias=False)
        self.bn = nn.BatchNorm2d(c2, eps=1e-3, momentum=0.03)
        self.act = nn.SiLU()

    def forward(self, x):
        return self.act(self.bn(self.conv(x)))


class Bottleneck(nn.Module):
    def __init__(self, c1, c2, shortcut=True, e=1.0):
        super().__init__()
        c_ = int(c2 * e)
        self.cv1 = Conv(c1, c_, 3, 1)
        self.cv2 = Conv(c_, c2, 3, 1)
        self.add = shortcut and c1 == c2

    def forward(self, x):
        return x + self.cv2(self.cv1(x)) if self.add else self.cv2(self.cv1(x))


class C2f(nn.Module):
    def __init__(self, c1, c2, n=1, shortcut=False, e=0.5):
        super().__init__()
        self.c = int(c2 * e)
        self.cv1 = Conv(c1, 2 * self.c, 1, 1)
        self.cv2 = Conv((2 + n) * self.c, c2, 1)
        self.m = nn.ModuleList(Bottleneck(self.c, self.c, shortcut, e=1.0) for _ in range(n))

    def forward(self, x):
        y = list(self.cv1(x).chunk(2, 1))
        y.extend(m(y[-1]) for m in self.m)
        return self.cv2(torch.cat(y, 1))


class SPPF(nn.Module):
    def __init__(self, c1, c2, k=5):
        super().__init__()
        c_ = c1 // 2
        self.cv1 = Conv(c1, c_, 1, 1)
        self.cv2 = Conv(c_ * 4, c2, 1, 1)
        self.m = nn.MaxPool2d(kernel_size=k, stride=1, padding=k // 2)

    def forward(self, x):
        x = self.cv1(x)
        y1 = self.m(x)
        y2 = self.m(y1)
        return self.cv2(torch.cat((x, y1, y2, self.m(y2)), 1))


def make_anchors(shapes, strides, offset=0.5):
    """Anchor centres (grid + 0.5) and stride per anchor, levels concatenated, row-major inside a level."""
    pts, strs = [], []
    for (h, w), s in zip(shapes, strides):
        sx = torch.arange(w, dtype=torch.float32) + offset
        sy = torch.arange(h, dtype=torch.float32) + offset
        yy, xx = torch.meshgrid(sy, sx, indexing="ij")
        pts.append(torch.stack((xx, yy), -1).view(-1, 2))
        strs.append(torch.full((h * w, 1), float(s)))
    return torch.cat(pts), torch.cat(strs)


class Detect(nn.Module):
    def __init__(self, nc=80, ch=(64, 128, 256)):
        super().__init__()
        self.nc, self.reg_max = nc, 16
        self.no = nc + 4 * self.reg_max
        self.stride = (8.0, 16.0, 32.0)
        c2 = max(16, ch[0] // 4, self.reg_max * 4)
        c3 = max(ch[0], min(nc, 100))
        self.cv2 = nn.ModuleList(nn.Sequential(Conv(x, c2, 3), Conv(c2, c2, 3), nn.Conv2d(c2, 4 * self.reg_max, 1)) for x in ch)
        self.cv3 = nn.ModuleList(nn.Sequential(Conv(x, c3, 3), Conv(c3, c3, 3), nn.Conv2d(c3, nc, 1)) for x in ch)

    def forward(self, feats):
        outs = [torch.cat((self.cv2[i](f), self.cv3[i](f)), 1) for i, f in enumerate(feats)]
        b = outs[0].shape[0]
        x_cat = torch.cat([o.view(b, self.no, -1) for o in outs], 2)
        box, cls = x_cat.split((4 * self.reg_max, self.nc), 1)
        anchors, strides = make_anchors([o.shape[2:] for o in outs], self.stride)
        # DFL: softmax over the 16 bins of each side, expectation
        a = box.shape[2]
        p = box.view(b, 4, self.reg_max, a).softmax(2)
        dist = (p * torch.arange(self.reg_max, dtype=torch.float32).view(1, 1, -1, 1)).sum(2)      # [b, 4, a] = l, t, r, b
        lt, rb = dist[:, :2], dist[:, 2:]
        anc = anchors.t().unsqueeze(0)                                                              # [1, 2, a]
        x1y1, x2y2 = anc - lt, anc + rb
        dbox = torch.cat(((x1y1 + x2y2) / 2, x2y2 - x1y1), 1) * strides.t().unsqueeze(0)            # xywh in pixels
        return torch.cat((dbox, cls.sigmoid()), 1)                                                  # [b, 4 + nc, a]


class YOLOv8n(nn.Module):
    """yolov8.yaml with the `n` scale: channels (16, 32, 64, 128, 256), C2f repeats (1, 2, 2, 1)."""

    def __init__(self, nc=80):
        super().__init__()
        self.m0 = Conv(3, 16, 3, 2)
        self.m1 = Conv(16, 32, 3, 2)
        self.m2 = C2f(32, 32, 1, True)
        self.m3 = Conv(32, 64, 3, 2)
        self.m4 = C2f(64, 64, 2, True)
        self.m5 = Conv(64, 128, 3, 2)
        self.m6 = C2f(128, 128, 2, True)
        self.m7 = Conv(128, 256, 3, 2)
        self.m8 = C2f(256, 256, 1, True)
        self.m9 = SPPF(256, 256, 5)
        self.m12 = C2f(384, 128, 1, False)
        self.m15 = C2f(192, 64, 1, False)
        self.m16 = Conv(64, 64, 3, 2)
        self.m18 = C2f(192, 128, 1, False)
        self.m19 = Conv(128, 128, 3, 2)
        self.m21 = C2f(384, 256, 1, False)
        self.m22 = Detect(nc, (64, 128, 256))

    def features(self, x):
        x = self.m2(self.m1(self.m0(x)))
        p3 = self.m4(self.m3(x))
        p4 = self.m6(self.m5(p3))
        p5 = self.m9(self.m8(self.m7(p4)))
        up = F.interpolate(p5, scale_factor=2.0, mode="nearest")
        h12 = self.m12(torch.cat((up, p4), 1))
        up = F.interpolate(h12, scale_factor=2.0, mode="nearest")
        h15 = self.m15(torch.cat((up, p3), 1))
        h18 = self.m18(torch.cat((self.m16(h15), h12), 1))
        h21 = self.m21(torch.cat((self.m19(h18), p5), 1))
        return h15, h18, h21

    def forward(self, x):
        """[B,3,640,640] -> [B, 84, 8400]: rows 0-3 xywh in input pixels, rows 4-83 class probabilities."""
        return self.m22(self.features(x))
