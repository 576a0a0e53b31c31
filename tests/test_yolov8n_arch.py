"""What can be pinned for YOLOv8n without ultralytics or its weights (neither is in the reference tree nor installable here):
the published parameter count, every yaml layer's output shape, the published FLOP count, and the agreement of two
separately written fp32 forwards -- the weight-carrying module (rdb_b200.yolov8n_arch) and the oracle's functional forward
written from the yaml table (oracle/yolov8n.py, which does not import the product).  Architecture pinned to the published
counts; numerics remain unpinned."""
import torch

import rdb_b200  # noqa: F401
from rdb_b200 import synth
from rdb_b200.yolov8n_arch import YOLOv8n
from oracle import yolov8n as oy


def test_parameter_count_matches_the_published_figure():
    m = YOLOv8n()
    n = sum(p.numel() for p in m.parameters())
    assert n == 3157184
    assert n + 16 == oy.PUBLISHED_PARAMS          # ultralytics counts the frozen DFL conv (16 weights), folded into the decode here


def test_flops_within_3_percent_of_the_published_figure():
    m = YOLOv8n().eval()
    flops, params = oy.conv_flops_and_params(m.state_dict())
    assert params == 3157184
    gf = flops / 1e9
    print("YOLOv8n conv FLOPs at 640x640: %.3f G (published %.1f)" % (gf, oy.PUBLISHED_GFLOPS))
    assert abs(gf - oy.PUBLISHED_GFLOPS) / oy.PUBLISHED_GFLOPS <= 0.03


def test_two_independent_forwards_agree_and_layer_shapes_match_the_yaml():
    m = synth.make_model("yolov8n").eval()
    x = synth.synthetic_images(2, 640, 640, seed=4242).permute(0, 3, 1, 2).float().div(255)
    with torch.no_grad():
        want = m(x)
        got, layers = oy.forward_functional(m.state_dict(), x, return_layers=True)
    assert want.shape == got.shape == (2, 84, 8400)
    assert (want - got).abs().max().item() <= 1e-3 * max(1.0, want.abs().max().item()) * 1e-2
    for idx, (c, stride) in oy.LAYER_SHAPES_640.items():
        assert tuple(layers[idx].shape) == (2, c, 640 // stride, 640 // stride), idx
    # anchors: 80 x 80 + 40 x 40 + 20 x 20, level by level, row-major
    assert 80 * 80 + 40 * 40 + 20 * 20 == 8400


def test_smaller_input_gives_proportionally_fewer_anchors():
    m = YOLOv8n().eval()
    x = torch.rand(1, 3, 320, 320)
    with torch.no_grad():
        assert tuple(oy.forward_functional(m.state_dict(), x).shape) == (1, 84, 2100)
        assert torch.allclose(m(x), oy.forward_functional(m.state_dict(), x), atol=1e-4, rtol=1e-4)
