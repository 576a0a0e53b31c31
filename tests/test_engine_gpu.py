"""End-to-end parity of the B200 engine (collate -> forward) through the C ABI, for every served model family.

Reference arm: the torchvision fp32 CPU forward the reference actor runs (scheduler.py:443-450), same synthetic weights.
Tolerances are the north-star ones: logits max-abs <= 2e-2, top-1 agreement >= 99.9 % on the synthetic set,
collate output bit-exact.
"""
import numpy as np
import pytest
import torch

import rdb_b200
from rdb_b200 import synth, lowering
from rdb_b200.engine import B200Model
from oracle import preprocess as opre
from oracle.plan_interp import PlanInterpreter, to_nhwc4

pytestmark = pytest.mark.gpu
LOGIT_TOL = 2e-2


KINDS = ["resnet50", "shufflenet_v2_x1_0", "efficientnet_v2_s"]


@pytest.fixture(scope="module", params=KINDS)
def resnet(request, cuda):
    """(torchvision fp32 module, B200 engine) for each model family; the name is historical."""
    kind = request.param
    module = synth.make_model(kind)
    model = B200Model(kind, module, max_batch=64)
    yield module, model
    model.close()


def _bits(host_bytes, shape):
    return host_bytes.numpy().view(np.uint16).reshape(shape)


def test_collate_identity_bit_exact(resnet):
    _, model = resnet
    imgs = synth.synthetic_images(5)
    packed, offsets, hw, total = model.pack_u8(list(imgs))
    model.collate_u8(packed, offsets, hw, 5, padded_batch=8, total_bytes=total)
    got = _bits(model.buffer(-1, 8), (8, 224, 224, 4))
    want = opre.collate_u8([im.numpy() for im in imgs], 8, 224, 224, synth.IMAGENET_MEAN, synth.IMAGENET_STD)
    assert np.array_equal(got, want)
    assert not got[5:].any(), "padding slots must be zero images"


def test_collate_f32_bit_exact(resnet):
    _, model = resnet
    reqs = synth.reference_requests(3)
    x = torch.stack(reqs)
    model.lib.b200_collate_f32(model._h, x.data_ptr(), 3, 4, 0, None, None)
    got = _bits(model.buffer(-1, 4), (4, 224, 224, 4))
    want = opre.collate_f32chw([r.numpy() for r in reqs], 4)
    assert np.array_equal(got, want)


def test_per_layer_against_plan_interpreter(resnet):
    """Every activation buffer of the CUDA engine vs the CPU interpreter of the same plan (same bf16 storage points)."""
    module, shared = resnet
    # an engine of its own that keeps every buffer: the serving engines let buffers with disjoint lifetimes share memory
    model = B200Model(shared.kind, blob=shared.blob, max_batch=4, keep_buffers=True)
    model.op_names = shared.op_names
    n = 3
    imgs = synth.synthetic_images(n)
    x = synth.normalize_u8(imgs, synth.IMAGENET_MEAN, synth.IMAGENET_STD)
    it = PlanInterpreter(model.blob, lowering.parse_blob)
    _, B = it.run(to_nhwc4(x), keep_all=True)
    logits = model.infer_u8(list(imgs))
    assert torch.equal(logits.cpu(), shared.infer_u8(list(imgs)).cpu()), "shared and unshared activation memory must agree"
    torch.cuda.synchronize()
    worst = []
    for op_i, op in enumerate(it.ops):
        buf = op["p"][3]
        h, w, c, dt = it.bufs[buf]
        raw = model.buffer(buf, n)
        if dt == 1:
            got = torch.from_numpy(raw.numpy().view(np.float32).reshape(n, h, w, c).copy())
        else:
            got = torch.from_numpy((raw.numpy().view(np.uint16).astype(np.uint32) << 16).view(np.float32)
                                   .reshape(n, h, w, c).copy())
        want = B[buf]
        err = (got - want).abs().max().item()
        scale = want.abs().max().item() + 1e-6
        worst.append((err / scale, op_i, model.op_names[op_i], err, scale))
        # storage rounding differences (1 bf16 ulp = 2^-8 relative) propagate and grow slowly through 50-230 layers
        assert err <= 0.06 * scale, "op %d (%s): max err %.4g vs scale %.4g" % (op_i, model.op_names[op_i], err, scale)
    worst.sort(reverse=True)
    print("worst layers:", worst[:3])
    keep_total, _ = model.memory_bytes()
    model.close()
    share_total, share_act = shared.memory_bytes()
    print("engine memory at max_batch %d: %.0f MB with shared activation memory" % (shared.max_batch, share_total / 2 ** 20))
    assert share_act > 0


def test_logits_and_top1_vs_fp32_cpu_reference(resnet):
    module, model = resnet
    n = synth.N_PROTO
    imgs = synth.synthetic_images(n)
    with torch.no_grad():
        ref = torch.cat([module(synth.normalize_u8(imgs[i:i + 32], synth.IMAGENET_MEAN, synth.IMAGENET_STD))
                         for i in range(0, n, 32)])
    got = torch.cat([model.infer_u8(list(imgs[i:i + 64])).cpu() for i in range(0, n, 64)])
    err = (got - ref).abs().max().item()
    rel = err / ref.std().item()
    agree = (got.argmax(1) == ref.argmax(1)).float().mean().item()
    planted = ref.diag().mean().item()
    print("%s: logits max-abs err %.4g (%.2f%% of the planted logit %.3g); top-1 agreement %.4f"
          % (model.kind, err, 100 * err / planted, planted, agree))
    assert err <= LOGIT_TOL
    assert agree >= 0.999
    assert (ref.argmax(1) == torch.arange(n)).all(), "reference itself must recover the planted classes"


def test_reference_call_site_randn_requests(resnet):
    """`outputs = model(inputs)` with the reference's own request shape: torch.randn(3,224,224) stacked (scheduler.py:443-450)."""
    module, model = resnet
    reqs = synth.reference_requests(9)
    inputs = torch.stack(reqs)
    with torch.no_grad():
        ref = module(inputs)
    got_dev = model(inputs.cuda()).cpu()
    got_host = model(inputs).cpu()
    assert torch.equal(got_dev, got_host)
    assert got_dev.shape == (9, 1000)
    assert (got_dev - ref).abs().max().item() <= LOGIT_TOL


def test_graph_replay_matches_eager_and_padding_rows_ignored(resnet):
    _, model = resnet
    imgs = list(synth.synthetic_images(6))
    a = model.infer_u8(imgs).cpu()                        # first call for batch 6: eager
    b = model.infer_u8(imgs).cpu()                        # second: captured graph
    c = model.infer_u8(imgs).cpu()                        # third: graph replay
    assert torch.equal(a, b) and torch.equal(b, c)
    d = model.infer_u8(imgs, padded_batch=8).cpu()        # padded batch: same logits for the real rows
    assert torch.equal(a, d)


def test_batch_size_independence(resnet):
    """Each request's logits do not depend on which batch it rides in (tile shapes change with batch size)."""
    _, model = resnet
    imgs = list(synth.synthetic_images(17))
    full = model.infer_u8(imgs).cpu()
    one = torch.cat([model.infer_u8(imgs[i:i + 1]).cpu() for i in (0, 5, 16)])
    assert torch.equal(full[[0, 5, 16]], one)


def test_engine_rejects_bad_input(resnet):
    _, model = resnet
    with pytest.raises(rdb_b200.lib.B200Error):
        model(torch.zeros(2, 3, 100, 100))
    with pytest.raises(rdb_b200.lib.B200Error):
        model(torch.zeros(65, 3, 224, 224))
    with pytest.raises(rdb_b200.lib.B200Error):
        B200Model("resnet50", blob=b"garbage" * 100)
