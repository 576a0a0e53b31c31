"""ORACLE (test infrastructure only — never imported by the product path).

YOLOv8n restated in plain PyTorch (fp32, CPU).  **Parity unpinned**: the reference ships only the image workload for this
model (`sample-models/yolov8/workload/*.JPEG`); its `yolov8n.pt/.onnx` blobs are missing (`.MISSING_LARGE_BLOBS`), no code in
the reference loads or runs it, and the `ultralytics` package (third-party, not vendored, not installed here; no pin in
the reference) cannot be imported.  No independent implementation exists to pin against; what CAN be pinned is pinned in tests/test_yolov8n_arch.py: the
parameter count (3,157,184 + the 16-element frozen DFL conv = the published 3,157,200), every yaml layer's output shape,
the FLOP count against the published 8.7 GFLOPs, and the agreement of two separately written forwards (the module in
rdb_b200.yolov8n_arch, which carries the weights, and forward_functional below, written from the yaml table and a state
dict; this file does not import the product).  Architecture pinned to the published counts; numerics unpinned.
This file restates the published architecture of ultralytics 8.0/8.1
(`cfg/models/v8/yolov8.yaml`, scale n = depth 0.33 / width 0.25; `nn/modules/conv.py` Conv, `nn/modules/block.py`
C2f / Bottleneck / SPPF / DFL, `nn/modules/head.py` Detect, `utils/tal.py` make_anchors / dist2bbox, `utils/ops.py`
non_max_suppression) from its public description; layer indices below are the yaml's.  IoU >= 0.99 is therefore judged
against this restatement with the same synthetic weights, not against ultralytics itself.
"""
import torch
import torch.nn.functional as F

# yolov8.yaml, scale n (depth 0.33 -> C2f repeats 1/2/2/1, width 0.25 -> channels 16..256): (index, kind, c_in, c_out, extra)
YAML_N = [
    (0, "conv", 3, 16, 2), (1, "conv", 16, 32, 2), (2, "c2f", 32, 32, (1, True)), (3, "conv", 32, 64, 2),
    (4, "c2f", 64, 64, (2, True)), (5, "conv", 64, 128, 2), (6, "c2f", 128, 128, (2, True)), (7, "conv", 128, 256, 2),
    (8, "c2f", 256, 256, (1, True)), (9, "sppf", 256, 256, 5),
    (10, "up", None, None, None), (11, "cat", (10, 6), 384, None), (12, "c2f", 384, 128, (1, False)),
    (13, "up", None, None, None), (14, "cat", (13, 4), 192, None), (15, "c2f", 192, 64, (1, False)),
    (16, "conv", 64, 64, 2), (17, "cat", (16, 12), 192, None), (18, "c2f", 192, 128, (1, False)),
    (19, "conv", 128, 128, 2), (20, "cat", (19, 9), 384, None), (21, "c2f", 384, 256, (1, False)),
]
# output (channels, stride) of every yaml layer at a 640 x 640 input: what tests/test_yolov8n_arch.py pins
LAYER_SHAPES_640 = {0: (16, 2), 1: (32, 4), 2: (32, 4), 3: (64, 8), 4: (64, 8), 5: (128, 16), 6: (128, 16), 7: (256, 32),
                    8: (256, 32), 9: (256, 32), 10: (256, 16), 11: (384, 16), 12: (128, 16), 13: (128, 8), 14: (192, 8),
                    15: (64, 8), 16: (64, 16), 17: (192, 16), 18: (128, 16), 19: (128, 32), 20: (384, 32), 21: (256, 32)}
PUBLISHED_PARAMS = 3157200       # ultralytics model summary for yolov8n (includes the frozen 16-element DFL conv)
PUBLISHED_GFLOPS = 8.7           # same summary, 640 x 640


def _cbs(sd, p, x, stride=1):
    """Conv2d(bias=False, pad k//2) + BatchNorm2d(eps 1e-3) + SiLU from state-dict entries `p`.conv / `p`.bn."""
    w = sd[p + ".conv.weight"]
    y = F.conv2d(x, w, None, stride, w.shape[2] // 2)
    y = F.batch_norm(y, sd[p + ".bn.running_mean"], sd[p + ".bn.running_var"], sd[p + ".bn.weight"], sd[p + ".bn.bias"],
                     False, 0.0, 1e-3)
    return y * torch.sigmoid(y)


def _c2f(sd, p, x, n, shortcut):
    y = _cbs(sd, p + ".cv1", x)
    half = y.shape[1] // 2
    parts = [y[:, :half], y[:, half:]]
    for i in range(n):
        t = _cbs(sd, "%s.m.%d.cv2" % (p, i), _cbs(sd, "%s.m.%d.cv1" % (p, i), parts[-1]))
        parts.append(parts[-1] + t if shortcut else t)
    return _cbs(sd, p + ".cv2", torch.cat(parts, 1))


def _sppf(sd, p, x, k):
    x = _cbs(sd, p + ".cv1", x)
    y1 = F.max_pool2d(x, k, 1, k // 2)
    y2 = F.max_pool2d(y1, k, 1, k // 2)
    y3 = F.max_pool2d(y2, k, 1, k // 2)
    return _cbs(sd, p + ".cv2", torch.cat((x, y1, y2, y3), 1))


def forward_functional(sd, x, prefix="m", return_layers=False):
    """YOLOv8n forward written a second time, straight from the yaml table above and a state dict whose keys follow the
    ultralytics naming (`<prefix><index>.cv1.conv.weight` ...): independent of any module class, so a slip in either
    restatement (concat order, C2f wiring, anchor order, DFL axis) shows up as a mismatch between the two.
    x: [B, 3, H, W] in [0, 1].  Returns [B, 84, A] (xywh in pixels, then 80 class probabilities)."""
    outs = {}
    for idx, kind, cin, cout, extra in YAML_N:
        p = "%s%d" % (prefix, idx)
        if kind == "conv":
            x = _cbs(sd, p, x, extra)
        elif kind == "c2f":
            x = _c2f(sd, p, x, *extra)
        elif kind == "sppf":
            x = _sppf(sd, p, x, extra)
        elif kind == "up":
            x = x.repeat_interleave(2, dim=2).repeat_interleave(2, dim=3)
        elif kind == "cat":
            x = torch.cat([outs[i] for i in cin], 1)
        outs[idx] = x
    p = "%s22" % prefix
    levels, strides = [outs[15], outs[18], outs[21]], (8.0, 16.0, 32.0)
    boxes, clss, centres, scale = [], [], [], []
    for i, f in enumerate(levels):
        b = _cbs(sd, "%s.cv2.%d.1" % (p, i), _cbs(sd, "%s.cv2.%d.0" % (p, i), f))
        b = F.conv2d(b, sd["%s.cv2.%d.2.weight" % (p, i)], sd["%s.cv2.%d.2.bias" % (p, i)])
        c = _cbs(sd, "%s.cv3.%d.1" % (p, i), _cbs(sd, "%s.cv3.%d.0" % (p, i), f))
        c = F.conv2d(c, sd["%s.cv3.%d.2.weight" % (p, i)], sd["%s.cv3.%d.2.bias" % (p, i)])
        n, _, h, w = b.shape
        boxes.append(b.reshape(n, 64, h * w))
        clss.append(c.reshape(n, 80, h * w))
        ys, xs = torch.meshgrid(torch.arange(h, dtype=torch.float32), torch.arange(w, dtype=torch.float32), indexing="ij")
        centres.append(torch.stack((xs.reshape(-1) + 0.5, ys.reshape(-1) + 0.5), 0))
        scale.append(torch.full((h * w,), strides[i]))
    box, cls = torch.cat(boxes, 2), torch.cat(clss, 2)
    ctr, st = torch.cat(centres, 1), torch.cat(scale)
    n, _, a = box.shape
    prob = torch.softmax(box.reshape(n, 4, 16, a), dim=2)                       # DFL: 16 bins per side
    dist = (prob * torch.arange(16, dtype=torch.float32).reshape(1, 1, 16, 1)).sum(2)   # l, t, r, b in grid units
    x1y1 = ctr.unsqueeze(0) - dist[:, :2]
    x2y2 = ctr.unsqueeze(0) + dist[:, 2:]
    xywh = torch.cat(((x1y1 + x2y2) * 0.5, x2y2 - x1y1), 1) * st
    out = torch.cat((xywh, torch.sigmoid(cls)), 1)
    return (out, outs) if return_layers else out


def conv_flops_and_params(sd):
    """2 x MAC of every convolution at 640 x 640 (from the weights' shapes and the yaml strides) and the parameter count."""
    params = sum(v.numel() for k, v in sd.items() if k.endswith((".weight", ".bias")))
    x = torch.zeros(1, 3, 640, 640)
    flops = [0.0]
    real = F.conv2d

    def counting(inp, w, b=None, stride=1, padding=0, *a, **k):
        y = real(inp, w, b, stride, padding, *a, **k)
        flops[0] += 2.0 * y.shape[2] * y.shape[3] * w.numel()
        return y
    F.conv2d = counting
    try:
        forward_functional(sd, x)
    finally:
        F.conv2d = real
    return flops[0], params


def xywh2xyxy(x):
    y = x.clone()
    y[..., 0] = x[..., 0] - x[..., 2] / 2
    y[..., 1] = x[..., 1] - x[..., 3] / 2
    y[..., 2] = x[..., 0] + x[..., 2] / 2
    y[..., 3] = x[..., 1] + x[..., 3] / 2
    return y


def box_iou(a, b):
    """IoU matrix of xyxy boxes [n,4] x [m,4]."""
    area_a = (a[:, 2] - a[:, 0]) * (a[:, 3] - a[:, 1])
    area_b = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
    lt = torch.max(a[:, None, :2], b[None, :, :2])
    rb = torch.min(a[:, None, 2:], b[None, :, 2:])
    inter = (rb - lt).clamp(min=0).prod(2)
    return inter / (area_a[:, None] + area_b[None] - inter + 1e-7)


def non_max_suppression(pred, conf_thres=0.25, iou_thres=0.45, max_det=300, max_wh=7680.0):
    """Greedy class-aware NMS on one image's [84, A] prediction (best class per anchor, class offset trick, score-descending;
    ties broken by anchor index).  Returns [k, 6]: x1, y1, x2, y2, score, class."""
    pred = pred.t()                                           # [A, 84]
    scores, cls = pred[:, 4:].max(1)
    keep = scores > conf_thres
    idx = torch.nonzero(keep).squeeze(1)
    if idx.numel() == 0:
        return torch.zeros(0, 6)
    boxes = xywh2xyxy(pred[idx, :4])
    sc, cl = scores[idx], cls[idx].float()
    order = torch.argsort(sc, descending=True, stable=True)
    boxes, sc, cl = boxes[order], sc[order], cl[order]
    shifted = boxes + (cl * max_wh).unsqueeze(1)
    kept = []
    for i in range(boxes.shape[0]):
        if len(kept) >= max_det:
            break
        if kept:
            if (box_iou(shifted[i:i + 1], shifted[torch.tensor(kept)]) > iou_thres).any():
                continue
        kept.append(i)
    k = torch.tensor(kept, dtype=torch.long)
    return torch.cat((boxes[k], sc[k, None], cl[k, None]), 1)
