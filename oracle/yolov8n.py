"""ORACLE (test infrastructure only — never imported by the product path).

YOLOv8n restated in plain PyTorch (fp32, CPU).  **Parity unpinned**: the reference ships only the image workload for this
model (`sample-models/yolov8/workload/*.JPEG`); its `yolov8n.pt/.onnx` blobs are missing (`.MISSING_LARGE_BLOBS`), no code in
the reference loads or runs it, and the `ultralytics` package (third-party, not vendored, not installed here; no pin in
the reference) cannot be imported.  This file restates the published architecture of ultralytics 8.0/8.1
(`cfg/models/v8/yolov8.yaml`, scale n = depth 0.33 / width 0.25; `nn/modules/conv.py` Conv, `nn/modules/block.py`
C2f / Bottleneck / SPPF / DFL, `nn/modules/head.py` Detect, `utils/tal.py` make_anchors / dist2bbox, `utils/ops.py`
non_max_suppression) from its public description; layer indices below are the yaml's.  IoU >= 0.99 is therefore judged
against this restatement with the same synthetic weights, not against ultralytics itself.
"""
import torch

import rdb_b200  # noqa: F401
from rdb_b200.yolov8n_arch import YOLOv8n, Detect, make_anchors  # noqa: F401  (the fp32 CPU forward is the reference arm)


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
