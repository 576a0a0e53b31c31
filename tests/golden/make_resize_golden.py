"""Generates tests/golden/resize_aa_golden.npz with TORCHVISION ITSELF (build container; torchvision 0.26 / torch 2.11):
torchvision.transforms.v2.functional.resize(uint8 CHW, antialias=True) -- the oracle SURVEY section 8 row P fixes for the
reference's transforms.Resize (profiling/run_profiler.py:63-70) -- on synthetic images whose SIZES are 22 of the 740 size
classes of the reference's sample-models/yolov8/workload (sizes only: the pixels are seeded noise + structure, so nothing
of the reference's data is copied).  Stored: the seeds, the sizes, and torchvision's outputs at 32 x 32 for every class
and at 112 x 112 for three (stretch and short-side + centre crop)."""
import os

import numpy as np
import torch
import torchvision.transforms.v2.functional as TF

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "resize_aa_golden.npz")
SIZES = [(314, 500), (513, 640), (500, 390), (240, 180), (390, 540), (540, 360), (480, 360), (576, 576), (500, 493),
         (175, 133), (540, 720), (102, 128), (479, 430), (326, 500), (500, 401), (312, 400), (48, 80), (3072, 2304),
         (500, 144), (203, 940), (224, 224), (500, 375)]          # (h, w)
FULL = [1, 17, 21]                                                  # classes also stored at 112 x 112


def make_image(i, h, w, gray=False):
    g = torch.Generator().manual_seed(77000 + i)
    c = 1 if gray else 3
    base = torch.nn.functional.interpolate(torch.rand(1, c, 9, 9, generator=g), size=(h, w), mode="bilinear")[0]
    img = ((base + (torch.rand(c, h, w, generator=g) - 0.5) * 0.5).clamp_(0, 1) * 255).round_().to(torch.uint8)
    return img                                                      # CHW uint8


def tv_resize(img_chw, oh, ow):
    return TF.resize(img_chw, [oh, ow], antialias=True).permute(1, 2, 0).numpy()


def tv_short_crop(img_chw, oh, ow):
    r = TF.resize(img_chw, [min(oh, ow)], antialias=True)
    return TF.center_crop(r, [oh, ow]).permute(1, 2, 0).numpy()


def main():
    out = {"sizes": np.asarray(SIZES, dtype=np.int32), "full": np.asarray(FULL, dtype=np.int32)}
    for i, (h, w) in enumerate(SIZES):
        img = make_image(i, h, w)
        out["stretch32_%d" % i] = tv_resize(img, 32, 32)
        out["short32_%d" % i] = tv_short_crop(img, 32, 32)
        if i in FULL:
            out["stretch112_%d" % i] = tv_resize(img, 112, 112)
            out["short112_%d" % i] = tv_short_crop(img, 112, 112)
    g = make_image(100, 368, 304, gray=True)                        # a grayscale request: PIL convert("RGB") replicates L
    out["gray_short32"] = tv_short_crop(g.expand(3, -1, -1).contiguous(), 32, 32)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes; torchvision", __import__("torchvision").__version__,
          "cpu capability", torch.backends.cpu.get_cpu_capability())


if __name__ == "__main__":
    main()
