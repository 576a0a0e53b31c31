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


def synthetic_images(n, h=224, w=224, seed=1000, dtype=torch.uint8):
    """Engine-native synthetic requests: uint8 HWC images, seed = 1000 + request index."""
    out = torch.empty(n, h, w, 3, dtype=torch.uint8)
    for i in range(n):
        g = torch.Generator().manual_seed(seed + i)
        out[i] = torch.randint(0, 256, (h, w, 3), generator=g, dtype=torch.uint8)
    return out


def reference_requests(n, h=224, w=224, seed=2000):
    """Reference-native requests: torch.randn(3,224,224) fp32 CHW (293-project/src/test_scheduler.py:89)."""
    return [torch.randn(3, h, w, generator=torch.Generator().manual_seed(seed + i)) for i in range(n)]


def _calibrate_bn(model, shape, seed, batches=2, batch_size=16):
    bns = [m for m in model.modules() if isinstance(m, nn.BatchNorm2d)]
    for m in bns:
        m.reset_running_stats()
        m.momentum = None  # cumulative moving average
    model.train()
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for _ in range(batches):
            model(torch.randn(batch_size, *shape, generator=g))
    model.eval()
    for m in bns:
        m.momentum = 0.1


def normalize_u8(imgs_u8, mean, std):
    """torchvision ToTensor + Normalize on a uint8 NHWC batch -> fp32 NCHW (profiling/run_profiler.py:63-70)."""
    x = imgs_u8.permute(0, 3, 1, 2).float().div(255)
    m = torch.tensor(mean, dtype=torch.float32).view(1, 3, 1, 1)
    s = torch.tensor(std, dtype=torch.float32).view(1, 3, 1, 1)
    return x.sub_(m).div_(s)


def _plant_classifier(model, fc: nn.Linear, in_hw, seed, n_proto, alpha, mean, std):
    """Nearest-prototype classifier: class j < n_proto fires on synthetic image j (seed 1000 + j); the remaining rows are
    random directions of the same scale.  logit_j(x) = alpha * <p_j, f(x) - mu>, p_j = (f_j - mu) / mean||f - mu||."""
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
    scale = d.norm(dim=1).mean()
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(fc.out_features, fc.in_features, generator=g, dtype=torch.float64)
    w = w / w.norm(dim=1, keepdim=True) * (d.norm(dim=1).mean() / scale)
    w[:n_proto] = d / scale
    with torch.no_grad():
        fc.weight.copy_((alpha * w).float())
        fc.bias.copy_((-alpha * (w @ mu)).float())


IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)
N_PROTO = 128  # classes 0..N_PROTO-1 are planted on synthetic images 1000..1000+N_PROTO-1


def make_resnet50(seed=0, damp=0.25, alpha=0.35, n_proto=N_PROTO):
    import torchvision
    torch.manual_seed(seed)
    m = torchvision.models.resnet50(weights=None)
    with torch.no_grad():
        for blk in [b for l in (m.layer1, m.layer2, m.layer3, m.layer4) for b in l]:
            blk.bn3.weight.fill_(damp)
    _calibrate_bn(m, (3, 224, 224), seed + 1)
    _plant_classifier(m, m.fc, (224, 224), seed + 2, n_proto, alpha, IMAGENET_MEAN, IMAGENET_STD)
    return m.eval()


MODEL_BUILDERS = {"resnet50": make_resnet50}


def make_model(kind, seed=0, **kw):
    if kind not in MODEL_BUILDERS:
        raise KeyError("no synthetic recipe for %r (have %s)" % (kind, sorted(MODEL_BUILDERS)))
    return MODEL_BUILDERS[kind](seed=seed, **kw)
