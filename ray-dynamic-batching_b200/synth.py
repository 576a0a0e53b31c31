"""Deterministic synthetic weights for the served model families (no network here: `pretrained=True` of
293-project/src/scheduler.py:40-44 cannot download).  The same module instance feeds both arms of every comparison:
the fp32 CPU reference path (torchvision forward) and the B200 engine (after lowering).

Recipe (seeded, so every machine builds bit-identical weights):
  1. torchvision's own initialisation under torch.manual_seed(seed);
  2. the last BatchNorm gamma of every residual branch is damped (like trained nets / zero-init-residual), which keeps
     activations O(1) through 16+ blocks instead of growing geometrically;
  3. BatchNorm running statistics are calibrated on seeded synthetic batches (cumulative average), so eval-mode
     activations are standardised like those of a trained network rather than degenerate;
  4. the classifier is re-drawn with a fixed gain so logits have O(1) spread (well-conditioned top-1 margins).
"""
import torch
import torch.nn as nn

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def synthetic_images(n, h=224, w=224, seed=1000, dtype=torch.uint8):
    """Engine-native synthetic requests: uint8 HWC images, one generator per request (seed + request index).
    Each image is a random coarse colour layout (7x7 control grid, bilinearly upsampled) plus pixel noise: unlike pure
    noise, different requests then differ in their low-frequency content, so the pooled features of different requests
    are well separated (what a classifier needs to have a top-1 at all)."""
    out = torch.empty(n, h, w, 3, dtype=torch.uint8)
    for i in range(n):
        g = torch.Generator().manual_seed(seed + i)
        grid = torch.rand(1, 3, 7, 7, generator=g)
        base = torch.nn.functional.interpolate(grid, size=(h, w), mode="bilinear", align_corners=False)[0]
        noise = (torch.rand(3, h, w, generator=g) - 0.5) * 0.25
        img = ((base + noise).clamp_(0, 1) * 255).round_()
        out[i] = img.permute(1, 2, 0).to(torch.uint8)
    return out


def heldout_images(n, h=224, w=224, seed=900000, jitter=0.08, n_proto=None):
    """Held-out requests for top-1 agreement: image i is a NEW sample of planted class i % n_proto -- the class's coarse
    colour layout (the 7x7 control grid of synthetic image 1000 + class) with every grid value jittered (sigma `jitter`)
    and fresh pixel noise, both from generator seed + i.  None of them is one of the images the classifier was planted
    on; the fp32 reference still recovers the class, with margins (top-1 minus top-2 logit) spread from a few
    hundredths up to ~0.5, so agreement on this set is decided by the engine's logit error, not by construction."""
    n_proto = n_proto or N_PROTO
    out = torch.empty(n, h, w, 3, dtype=torch.uint8)
    for i in range(n):
        g0 = torch.Generator().manual_seed(1000 + i % n_proto)
        grid = torch.rand(1, 3, 7, 7, generator=g0)
        g = torch.Generator().manual_seed(seed + i)
        grid = (grid + jitter * torch.randn(1, 3, 7, 7, generator=g)).clamp_(0, 1)
        base = torch.nn.functional.interpolate(grid, size=(h, w), mode="bilinear", align_corners=False)[0]
        noise = (torch.rand(3, h, w, generator=g) - 0.5) * 0.25
        img = ((base + noise).clamp_(0, 1) * 255).round_()
        out[i] = img.permute(1, 2, 0).to(torch.uint8)
    return out


def reference_requests(n, h=224, w=224, seed=2000):
    """Reference-native requests: torch.randn(3,224,224) fp32 CHW (293-project/src/test_scheduler.py:89)."""
    return [torch.randn(3, h, w, generator=torch.Generator().manual_seed(seed + i)) for i in range(n)]


def _calibrate_bn(model, shape, seed, batches=2, batch_size=16):
    """BatchNorm running statistics from seeded synthetic requests (half structured images, half randn tensors — the two
    request kinds the engine serves), cumulative average."""
    bns = [m for m in model.modules() if isinstance(m, nn.BatchNorm2d)]
    for m in bns:
        m.reset_running_stats()
        m.momentum = None  # cumulative moving average
    model.train()
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for b in range(batches):
            half = batch_size // 2
            imgs = synthetic_images(half, shape[1], shape[2], seed=500000 + seed * 1000 + b * half)
            x = torch.cat([normalize_u8(imgs, IMAGENET_MEAN, IMAGENET_STD),
                           torch.randn(batch_size - half, *shape, generator=g)])
            model(x)
    model.eval()
    for m in bns:
        m.momentum = 0.1


def normalize_u8(imgs_u8, mean, std):
    """torchvision ToTensor + Normalize on a uint8 NHWC batch -> fp32 NCHW (profiling/run_profiler.py:63-70)."""
    x = imgs_u8.permute(0, 3, 1, 2).float().div(255)
    m = torch.tensor(mean, dtype=torch.float32).view(1, 3, 1, 1)
    s = torch.tensor(std, dtype=torch.float32).view(1, 3, 1, 1)
    return x.sub_(m).div_(s)


def _plant_classifier(model, fc: nn.Linear, in_hw, seed, n_proto, planted, mean, std, ridge=0.05):
    """Ridge-whitened nearest-prototype classifier: class j < n_proto fires on synthetic image j (seed 1000 + j) with a
    logit of about `planted` and stays near 0 on the other prototype images; the remaining rows are random directions
    of the same norm.  With D the centred pooled features of the prototype images (fp32 reference forward),
    W[:n_proto] = planted' * (D D^T + ridge * mean diag)^-1 D,  b = -W mu."""
    feats = []
    hook = fc.register_forward_pre_hook(lambda mod, inp: feats.append(inp[0].detach().clone()))
    imgs = synthetic_images(n_proto, in_hw[0], in_hw[1])
    with torch.no_grad():
        for i in range(0, n_proto, 32):
            model(normalize_u8(imgs[i:i + 32], mean, std))
    hook.remove()
    f = torch.cat(feats).double()
    mu = f.mean(0)
    d = f - mu
    gram = d @ d.t()
    lam = ridge * gram.diag().mean()
    wp = torch.linalg.solve(gram + lam * torch.eye(n_proto, dtype=torch.float64), d)
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(fc.out_features, fc.in_features, generator=g, dtype=torch.float64)
    w = w / w.norm(dim=1, keepdim=True) * wp.norm(dim=1).mean()
    w[:n_proto] = wp
    gain = planted / float((wp * d).sum(1).mean())
    with torch.no_grad():
        fc.weight.copy_((gain * w).float())
        fc.bias.copy_((-gain * (w @ mu)).float())


def _plant_classifier_lda(model, fc: nn.Linear, in_hw, seed, n_proto, planted, mean, std, samples=4, jitter=0.08, lam=1.0):
    """Linear-discriminant head: `samples` jittered samples of each of the n_proto classes (heldout_images with seeds
    600000 + 1000 s: disjoint from the evaluation seeds) give class means mu_c and the within-class scatter S_w of the
    pooled fp32 features; W[:n_proto] = planted' * (S_w + lam * mean diag)^-1 (mu_c - mu), b = -W mu, remaining rows random
    directions of the same norm.  Unlike the nearest-prototype head this one does not key on one noise realisation: the
    fp32 reference classifies NEW samples of a class correctly, which is what a top-1 agreement figure needs."""
    feats = []
    hook = fc.register_forward_pre_hook(lambda mod, inp: feats.append(inp[0].detach().clone()))
    with torch.no_grad():
        for s_i in range(samples):
            imgs = heldout_images(n_proto, in_hw[0], in_hw[1], seed=600000 + 1000 * s_i, jitter=jitter, n_proto=n_proto)
            for i in range(0, n_proto, 32):
                model(normalize_u8(imgs[i:i + 32], mean, std))
    hook.remove()
    f = torch.cat(feats).double().view(samples, n_proto, -1)
    mu_c = f.mean(0)
    mu = mu_c.mean(0)
    dw = (f - mu_c).reshape(samples * n_proto, -1)
    sw = dw.t() @ dw / (samples * n_proto)
    d = mu_c - mu
    a = sw + lam * sw.diag().mean() * torch.eye(sw.shape[0], dtype=torch.float64)
    wp = torch.linalg.solve(a, d.t()).t()
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(fc.out_features, fc.in_features, generator=g, dtype=torch.float64)
    w = w / w.norm(dim=1, keepdim=True) * wp.norm(dim=1).mean()
    w[:n_proto] = wp
    gain = planted / float((wp * d).sum(1).mean())
    with torch.no_grad():
        fc.weight.copy_((gain * w).float())
        fc.bias.copy_((-gain * (w @ mu)).float())


def _plant(head, *a, **k):
    if head == "prototype":
        return _plant_classifier(*a, **k)
    if head == "lda":
        return _plant_classifier_lda(*a, **k)
    raise ValueError("head must be 'prototype' or 'lda'")


N_PROTO = 128  # classes 0..N_PROTO-1 are planted on synthetic images 1000..1000+N_PROTO-1


def make_resnet50(seed=0, damp=0.25, planted=1.5, n_proto=N_PROTO, head="prototype"):
    import torchvision
    torch.manual_seed(seed)
    m = torchvision.models.resnet50(weights=None)
    with torch.no_grad():
        for blk in [b for l in (m.layer1, m.layer2, m.layer3, m.layer4) for b in l]:
            blk.bn3.weight.fill_(damp)
    _calibrate_bn(m, (3, 224, 224), seed + 1)
    _plant(head, m, m.fc, (224, 224), seed + 2, n_proto, planted, IMAGENET_MEAN, IMAGENET_STD)
    return m.eval()


def make_shufflenet(seed=0, planted=0.8, n_proto=N_PROTO, bn_bias=1.0, head="prototype"):
    """No residual adds to damp here; a positive BatchNorm beta keeps most ReLU inputs positive, i.e. the random network
    out of the chaotic regime where a 2^-9 rounding perturbation grows ~1.4x per unit (measured with the CPU interpreter)."""
    import torchvision
    torch.manual_seed(seed)
    m = torchvision.models.shufflenet_v2_x1_0(weights=None)
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, nn.BatchNorm2d):
                mod.bias.fill_(bn_bias)
    _calibrate_bn(m, (3, 224, 224), seed + 1)
    _plant(head, m, m.fc, (224, 224), seed + 2, n_proto, planted, IMAGENET_MEAN, IMAGENET_STD)
    return m.eval()


def make_efficientnet_v2_s(seed=0, damp=0.25, planted=1.0, n_proto=N_PROTO, in_hw=(224, 224), head="prototype"):
    import torchvision
    torch.manual_seed(seed)
    m = torchvision.models.efficientnet_v2_s(weights=None)
    with torch.no_grad():
        for stage in m.features:
            for blk in stage:
                if getattr(blk, "use_res_connect", False):
                    blk.block[-1][1].weight.fill_(damp)     # BatchNorm of the block's last conv (residual branch)
    _calibrate_bn(m, (3, in_hw[0], in_hw[1]), seed + 1, batches=2, batch_size=8)
    _plant(head, m, m.classifier[1], in_hw, seed + 2, n_proto, planted, IMAGENET_MEAN, IMAGENET_STD)
    return m.eval()


def make_yolov8n(seed=0, damp=0.25, cls_bias=-6.0, cls_gain=3.0, box_gain=0.5, bn_bias=1.0, in_hw=(640, 640)):
    """Seeded YOLOv8n: default PyTorch init, last BN of every shortcut bottleneck damped, BN statistics calibrated on seeded
    640x640 requests (scaled to [0,1], no mean/std: YOLO preprocessing), Detect biases set so that only a small fraction
    of anchors clears the 0.25 confidence threshold (a usable NMS workload)."""
    from .yolov8n_arch import YOLOv8n, Bottleneck
    torch.manual_seed(seed)
    m = YOLOv8n()
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, nn.BatchNorm2d):
                mod.bias.fill_(bn_bias)            # keeps the random net out of the chaotic regime (see make_shufflenet)
        for mod in m.modules():
            if isinstance(mod, Bottleneck) and mod.add:
                mod.cv2.bn.weight.fill_(damp)
        for i in range(3):
            m.m22.cv2[i][2].bias.fill_(1.0)
            m.m22.cv2[i][2].weight.mul_(box_gain)
            m.m22.cv3[i][2].bias.fill_(cls_bias)
            m.m22.cv3[i][2].weight.mul_(cls_gain)
    bns = [x for x in m.modules() if isinstance(x, nn.BatchNorm2d)]
    for b in bns:
        b.reset_running_stats()
        b.momentum = None
    m.train()
    with torch.no_grad():
        for bi in range(2):
            imgs = synthetic_images(4, in_hw[0], in_hw[1], seed=700000 + seed * 100 + bi * 4)
            m.features(imgs.permute(0, 3, 1, 2).float().div(255))
    m.eval()
    for b in bns:
        b.momentum = 0.03
    return m


MODEL_BUILDERS = {"yolov8n": make_yolov8n, "resnet50": make_resnet50, "shufflenet_v2_x1_0": make_shufflenet,
                  "efficientnet_v2_s": make_efficientnet_v2_s}


def make_model(kind, seed=0, **kw):
    if kind not in MODEL_BUILDERS:
        raise KeyError("no synthetic recipe for %r (have %s)" % (kind, sorted(MODEL_BUILDERS)))
    return MODEL_BUILDERS[kind](seed=seed, **kw)
