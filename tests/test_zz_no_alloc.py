"""b200_collate_u8_noalloc — the collate entry point for callers that may not allocate device memory mid-flight (a model
actor packed onto a GPU by its profiled peak memory: no headroom for, and no tolerance of the device-wide stall of, a
cudaMalloc inside its duty cycle).  It must produce exactly what the regular b200_collate produces; the collate tests
are bit-exact, so this one is too.  Two calls are made back to back on purpose: all calls share one fixed scratch array
inside the library, and the second call has to work on top of whatever the first one left there.

Collected last (file name sorts after every other test file) so a `-x` run has already exercised everything else.
"""
import ctypes

import numpy as np
import pytest
import torch

import rdb_b200
from rdb_b200 import lowering, synth
from rdb_b200.engine import B200Model, COLLATE_IDENTITY, COLLATE_LETTERBOX
from rdb_b200.lib import check, ptr, stream_ptr
from oracle import preprocess as opre

pytestmark = pytest.mark.gpu
MEAN, STD, FILL = synth.IMAGENET_MEAN, synth.IMAGENET_STD, 114
MAX_IMAGES = 1024  # B200_NOALLOC_MAX_IMAGES in include/b200infer.h


def _tiny_engine(h, w):
    """An engine with no ops: only its input buffer and the regular b200_collate are used."""
    pb = lowering.PlanBuilder("tiny", h, w, mean=MEAN, std=STD, pre_fill=FILL)
    pb.set_output(lowering.T(pb.new_buf(1, 1, 8, dtype=1), 0, 8, 1, 1))
    return B200Model("tiny", blob=pb.to_blob(), max_batch=8)


def _pack(images, device):
    offsets, hw, chunks, total = [], [], [], 0
    for im in images:
        offsets.append(total)
        hw.append((im.shape[0], im.shape[1]))
        flat = im.reshape(-1)
        pad = (-flat.numel()) % 16
        chunks.append(torch.cat([flat, torch.zeros(pad, dtype=torch.uint8)]))
        total += flat.numel() + pad
    packed = torch.cat(chunks).to(device)
    return packed, np.asarray(offsets, dtype=np.uint64), np.asarray(hw, dtype=np.int32), total


def _existing(model, packed, offsets, hw, n, padded, mode, total):
    model.collate_u8(packed, offsets, hw, n, padded, mode, total)
    return model.buffer(-1, padded).numpy().view(np.uint16).copy()


def _noalloc(lib, packed, offsets, hw, n, padded, h, w, mode, total, device):
    out = torch.full((padded, h, w, 4), -1, dtype=torch.int16, device=device)      # poisoned: every element must be written
    mean = (ctypes.c_float * 3)(*MEAN)
    std = (ctypes.c_float * 3)(*STD)
    check(lib.b200_collate_u8_noalloc(ptr(packed), offsets.ctypes.data_as(ctypes.c_void_p),
                                      hw.ctypes.data_as(ctypes.c_void_p), n, padded, h, w, mode, FILL, mean, std, total,
                                      ptr(out), stream_ptr()), "b200_collate_u8_noalloc")
    torch.cuda.synchronize()
    return out.cpu().numpy().view(np.uint16).reshape(padded, -1)


def test_noalloc_collate_matches_regular_collate_on_two_consecutive_calls(b200lib, cuda):
    model = _tiny_engine(224, 224)
    try:
        # call 1: five same-size requests padded to a batch of 8 (vectorised flat kernel)
        imgs1 = list(synth.synthetic_images(5, seed=4000))
        p1, o1, hw1, t1 = _pack(imgs1, cuda)
        want1 = _existing(model, p1, o1, hw1, 5, 8, COLLATE_IDENTITY, t1)
        got1 = _noalloc(b200lib, p1, o1, hw1, 5, 8, 224, 224, COLLATE_IDENTITY, t1, cuda)
        assert np.array_equal(got1, want1)
        # ... and against the CPU restatement of ToTensor + Normalize + stack, not only against the sibling entry point
        oracle1 = opre.collate_u8([im.numpy() for im in imgs1], 8, 224, 224, MEAN, STD)
        assert np.array_equal(got1.reshape(8, 224, 224, 4), oracle1)
        # call 2, straight after: three ragged requests, letterboxed (general kernel), different table contents
        g = torch.Generator().manual_seed(5)
        imgs2 = [torch.randint(0, 256, s, generator=g, dtype=torch.uint8) for s in ((75, 333, 3), (480, 640, 3), (224, 224, 3))]
        p2, o2, hw2, t2 = _pack(imgs2, cuda)
        want2 = _existing(model, p2, o2, hw2, 3, 3, COLLATE_LETTERBOX, t2)
        got2 = _noalloc(b200lib, p2, o2, hw2, 3, 3, 224, 224, COLLATE_LETTERBOX, t2, cuda)
        assert np.array_equal(got2, want2)
        oracle2 = opre.collate_u8([im.numpy() for im in imgs2], 3, 224, 224, MEAN, STD, mode=opre.MODE_LETTERBOX, fill=FILL)
        assert np.array_equal(got2.reshape(3, 224, 224, 4), oracle2)
        assert not np.array_equal(got2[:3], got1[:3])
    finally:
        model.close()


def test_noalloc_collate_refuses_more_images_than_its_fixed_scratch(b200lib, cuda):
    n = MAX_IMAGES + 1
    dummy = torch.zeros(64, dtype=torch.uint8, device=cuda)
    offsets = np.zeros(n, dtype=np.uint64)
    hw = np.ones((n, 2), dtype=np.int32)
    mean, std = (ctypes.c_float * 3)(*MEAN), (ctypes.c_float * 3)(*STD)
    rc = b200lib.b200_collate_u8_noalloc(ptr(dummy), offsets.ctypes.data_as(ctypes.c_void_p),
                                         hw.ctypes.data_as(ctypes.c_void_p), n, n, 1, 1, 1, 0, mean, std, 64, ptr(dummy),
                                         None)
    assert rc == -1
    msg = b200lib.b200_last_error().decode()
    assert "exceed" in msg and str(MAX_IMAGES) in msg


def test_noalloc_collate_leaves_device_memory_unchanged_after_warm_up_and_takes_unaligned_input(b200lib, cuda):
    """After one warm-up call on the device, a call changes neither the free-memory figure of the driver nor torch's
    allocator state; a packed buffer that is not 16-byte aligned is served too (byte-wise kernel), with the same result."""
    imgs = list(synth.synthetic_images(3, seed=4100))
    p, o, hw, t = _pack(imgs, cuda)
    want = opre.collate_u8([im.numpy() for im in imgs], 4, 224, 224, MEAN, STD)
    _noalloc(b200lib, p, o, hw, 3, 4, 224, 224, COLLATE_IDENTITY, t, cuda)          # warm-up (address lookup, code load)
    shifted = torch.empty(t + 16, dtype=torch.uint8, device=cuda)
    shifted[1:1 + t].copy_(p[:t])
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    got = _noalloc(b200lib, p, o, hw, 3, 4, 224, 224, COLLATE_IDENTITY, t, cuda)
    got_shifted = _noalloc(b200lib, shifted[1:], o, hw, 3, 4, 224, 224, COLLATE_IDENTITY, t, cuda)
    free1, _ = torch.cuda.mem_get_info()
    # _noalloc itself allocates the output through torch's caching allocator; the driver-level figure may only move by
    # what torch reserved for those outputs (it reuses its cached blocks here), never by the entry point
    assert free1 >= free0 - 2 * got.nbytes
    assert np.array_equal(got.reshape(4, 224, 224, 4), want)
    assert np.array_equal(got_shifted.reshape(4, 224, 224, 4), want)


def test_noalloc_collate_refuses_a_capturing_stream(b200lib, cuda):
    imgs = list(synth.synthetic_images(1, seed=4200))
    p, o, hw, t = _pack(imgs, cuda)
    out = torch.empty(1, 224, 224, 4, dtype=torch.int16, device=cuda)
    mean, std = (ctypes.c_float * 3)(*MEAN), (ctypes.c_float * 3)(*STD)
    _noalloc(b200lib, p, o, hw, 1, 1, 224, 224, COLLATE_IDENTITY, t, cuda)          # warm-up outside capture
    s = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    rc = None
    with torch.cuda.stream(s):
        g.capture_begin()
        try:
            rc = b200lib.b200_collate_u8_noalloc(ptr(p), o.ctypes.data_as(ctypes.c_void_p), hw.ctypes.data_as(ctypes.c_void_p),
                                                 1, 1, 224, 224, COLLATE_IDENTITY, FILL, mean, std, t, ptr(out),
                                                 ctypes.c_void_p(s.cuda_stream))
        finally:
            g.capture_end()
    assert rc == -1 and "captur" in b200lib.b200_last_error().decode()
