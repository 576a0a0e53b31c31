"""YOLOv8n through the C ABI: letterbox collate, forward, fused head decode, NMS.

Reference arm: the plain-PyTorch fp32 restatement (rdb_b200.yolov8n_arch / oracle/yolov8n.py; parity unpinned by the
reference, which never runs this model) with the same seeded weights.  Bars: boxes IoU >= 0.99 for every anchor
(north star), class probabilities within 2e-2, NMS identical to the oracle NMS on the same head tensor."""
import numpy as np
import pytest
import torch

import rdb_b200
from rdb_b200 import lowering, synth
from rdb_b200.engine import B200Model, COLLATE_LETTERBOX
from oracle import preprocess as opre
from oracle.plan_interp import PlanInterpreter, to_nhwc4
from oracle.yolov8n import box_iou, non_max_suppression, xywh2xyxy

pytestmark = pytest.mark.gpu
A = 8400


@pytest.fixture(scope="module")
def yolo(cuda):
    module = synth.make_model("yolov8n")
    model = B200Model("yolov8n", module, max_batch=8)
    yield module, model
    model.close()


def pair_iou(a, b):
    lt = torch.max(a[..., :2], b[..., :2])
    rb = torch.min(a[..., 2:], b[..., 2:])
    inter = (rb - lt).clamp(min=0).prod(-1)
    return inter / ((a[..., 2] - a[..., 0]) * (a[..., 3] - a[..., 1]) + (b[..., 2] - b[..., 0]) * (b[..., 3] - b[..., 1]) - inter)


def test_head_boxes_and_scores_vs_fp32_reference(yolo):
    module, model = yolo
    n = 3
    imgs = synth.synthetic_images(n, 640, 640)
    with torch.no_grad():
        ref = module(imgs.permute(0, 3, 1, 2).float().div(255))                  # [n, 84, 8400]
    got = model.infer_u8(list(imgs)).cpu().view(n, 84, A)
    iou = pair_iou(xywh2xyxy(ref[:, :4].permute(0, 2, 1)), xywh2xyxy(got[:, :4].permute(0, 2, 1)))
    cls_err = (got[:, 4:] - ref[:, 4:]).abs().max().item()
    print("yolov8n: min box IoU %.5f over %d anchors, max class-prob err %.4g" % (iou.min().item(), iou.numel(), cls_err))
    assert iou.min().item() >= 0.99
    assert cls_err <= 2e-2
    assert model.out_dim == 84 * A


def test_per_layer_against_plan_interpreter(yolo):
    module, shared = yolo
    model = B200Model("yolov8n", blob=shared.blob, max_batch=2, keep_buffers=True)   # every buffer readable afterwards
    model.op_names = shared.op_names
    n = 2
    imgs = synth.synthetic_images(n, 640, 640)
    x = imgs.permute(0, 3, 1, 2).float().div(255)
    it = PlanInterpreter(model.blob, lowering.parse_blob)
    _, B = it.run(to_nhwc4(x), keep_all=True)
    out = model.infer_u8(list(imgs))
    assert torch.equal(out.cpu(), shared.infer_u8(list(imgs)).cpu()), "shared and unshared activation memory must agree"
    torch.cuda.synchronize()
    for op_i, op in enumerate(it.ops):
        if op["type"] == lowering.OP_YOLO_DECODE:
            continue
        buf = op["p"][3]
        h, w, c, dt = it.bufs[buf]
        raw = model.buffer(buf, n)
        got = torch.from_numpy((raw.numpy().view(np.uint16).astype(np.uint32) << 16).view(np.float32).reshape(n, h, w, c).copy())
        want = B[buf]
        err = (got - want).abs().max().item()
        scale = want.abs().max().item() + 1e-6
        assert err <= 0.06 * scale, "op %d (%s): max err %.4g vs scale %.4g" % (op_i, model.op_names[op_i], err, scale)


def test_nms_kernel_matches_oracle_on_the_same_head(yolo):
    module, model = yolo
    n = 4
    imgs = synth.synthetic_images(n, 640, 640, seed=3000)
    head = model.infer_u8(list(imgs))
    boxes, counts = model.decode_yolo(n)
    torch.cuda.synchronize()
    head_cpu = head.cpu().view(n, 84, A)
    total = 0
    for i in range(n):
        want = non_max_suppression(head_cpu[i])
        k = int(counts[i].item())
        got = boxes[i, :k].cpu()
        assert k == want.shape[0], "image %d: %d vs %d detections" % (i, k, want.shape[0])
        assert torch.allclose(got[:, :5], want[:, :5], atol=1e-3, rtol=1e-5)
        assert torch.equal(got[:, 5], want[:, 5])
        total += k
    assert total > 0, "the synthetic workload must produce detections"


def test_detections_vs_fp32_reference_after_nms(yolo):
    """Confident reference detections (score well above the threshold) are reproduced with IoU >= 0.99 and the same class."""
    module, model = yolo
    n = 2
    imgs = synth.synthetic_images(n, 640, 640, seed=3100)
    with torch.no_grad():
        ref = module(imgs.permute(0, 3, 1, 2).float().div(255))
    model.infer_u8(list(imgs))
    boxes, counts = model.decode_yolo(n, conf=0.2)
    for i in range(n):
        want = non_max_suppression(ref[i], conf_thres=0.3)
        got = boxes[i, :int(counts[i].item())].cpu()
        if want.shape[0] == 0:
            continue
        iou = box_iou(want[:, :4], got[:, :4])
        best, arg = iou.max(1)
        hit = (best >= 0.99) & (got[arg, 5] == want[:, 5])
        assert hit.float().mean().item() >= 0.9, "only %.2f of the confident reference detections were reproduced" % hit.float().mean().item()


def test_letterbox_collate_of_ragged_requests_bit_exact(yolo):
    _, model = yolo
    g = torch.Generator().manual_seed(11)
    sizes = [(480, 640), (75, 333), (640, 640), (1000, 700), (31, 31)]
    imgs = [torch.randint(0, 256, (h, w, 3), generator=g, dtype=torch.uint8) for h, w in sizes]
    packed, offsets, hw, total = model.pack_u8(imgs)
    model.collate_u8(packed, offsets, hw, len(imgs), padded_batch=6, mode=COLLATE_LETTERBOX, total_bytes=total)
    got = model.buffer(-1, 6).numpy().view(np.uint16).reshape(6, 640, 640, 4)
    want = opre.collate_u8([im.numpy() for im in imgs], 6, 640, 640, (0, 0, 0), (1, 1, 1), mode=opre.MODE_LETTERBOX, fill=114)
    assert np.array_equal(got, want)
