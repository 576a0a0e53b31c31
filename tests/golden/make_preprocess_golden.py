"""Generates tests/golden/preprocess_golden.npz from torchvision's ToTensor + Normalize (the reference's client-side
preprocessing, profiling/run_profiler.py:63-70) followed by torch.stack (scheduler.py:443), on two of the reference's
own dataset images (293-project/dataset/cat*.jpg, decoded with PIL) plus one seeded random image. Run in the build
container only; the .npz is what travels."""
import os

import numpy as np
import torch
from PIL import Image
from torchvision import transforms

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "preprocess_golden.npz")
MEAN, STD = [0.485, 0.456, 0.406], [0.229, 0.224, 0.225]


def main():
    imgs = []
    for name in ("cat1.jpg", "cat2.jpg"):
        p = os.path.join("/root/reference/293-project/dataset", name)
        imgs.append(np.asarray(Image.open(p).convert("RGB")))
    imgs.append(np.random.default_rng(0).integers(0, 256, (224, 224, 3), dtype=np.uint8))
    assert all(im.shape == (224, 224, 3) for im in imgs)
    t = transforms.Compose([transforms.ToTensor(), transforms.Normalize(mean=MEAN, std=STD)])
    stacked = torch.stack([t(Image.fromarray(im)) for im in imgs])          # [3,3,224,224] fp32
    padded = 4
    bits = np.zeros((padded, 224, 224, 4), dtype=np.uint16)
    bits[:3, :, :, :3] = stacked.to(torch.bfloat16).view(torch.int16).numpy().astype(np.uint16).transpose(0, 2, 3, 1)
    np.savez_compressed(OUT, images=np.stack(imgs), expected_bits=bits, padded_batch=padded)
    print("wrote", OUT, os.path.getsize(OUT))


if __name__ == "__main__":
    main()
