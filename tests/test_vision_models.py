"""paddle.vision.models: every family builds, has the published parameter count, names its sub-layers as the
reference does (so .pdparams checkpoints load by key), runs forward / backward (reference:
python/paddle/vision/models/*.py)."""
import re

import pytest
import torch

import paddle
from paddle.vision import models


def _count(m):
    return sum(p.numel() for p in m.parameters())


def test_reference_export_list_is_complete():
    src = open("/root/reference/python/paddle/vision/models/__init__.py").read() if __import__("os").path.exists(
        "/root/reference/python/paddle/vision/models/__init__.py") else None
    if src is None:
        pytest.skip("reference tree not available")
    names = re.findall(r"'([^']+)'", re.search(r"__all__ = \[(.*?)\]", src, re.S).group(1))
    assert [n for n in names if not hasattr(models, n)] == []


# published ImageNet-1k parameter counts (BatchNorm running statistics are buffers, not counted)
@pytest.mark.parametrize("name,count", [
    ("mobilenet_v1", 4231976), ("mobilenet_v3_small", 2542856), ("mobilenet_v3_large", 5483032), ("densenet121", 7978856),
    ("squeezenet1_0", 1248424), ("squeezenet1_1", 1235496), ("shufflenet_v2_x0_5", 1366792), ("shufflenet_v2_x1_0", 2278604),
    ("inception_v3", 23834568), ("resnext50_32x4d", 25028904), ("wide_resnet50_2", 68883240), ("resnet50", 25557032),
])
def test_parameter_counts(name, count):
    assert _count(getattr(models, name)()) == count


def test_forward_backward_and_layer_names():
    x = torch.randn(2, 3, 64, 64)
    cases = {
        "mobilenet_v1": ["conv1.0.weight", "conv2_1._depthwise_conv.0.weight", "conv5_6._pointwise_conv.1.weight", "conv6._pointwise_conv.1._mean", "fc.weight"],
        "mobilenet_v3_small": ["conv.0.weight", "blocks.0.bottleneck_conv.0.weight", "blocks.0.mid_se.fc1.bias", "blocks.1.expand_conv.1.weight",
                               "lastconv.1._variance", "classifier.0.weight", "classifier.3.bias"],
        "densenet121": ["conv1_func._conv.weight", "db_conv_2.conv2_1.bn_ac_func1._batch_norm.weight", "db_conv_5.conv5_16.bn_ac_func2._conv.weight",
                        "tr_conv3_blk.conv_ac_func._conv.weight", "batch_norm._mean", "out.bias"],
        "squeezenet1_1": ["_conv.weight", "_conv1._conv._conv.weight", "_conv8._conv_path2._conv.bias", "_conv9.weight"],
        "shufflenet_v2_x0_25": ["_conv1.0.weight", "2_1._conv_dw_1.0.weight", "2_2._conv_pw.1.weight", "4_4._conv_linear.0.weight", "_last_conv.0.weight", "_fc.bias"],
        "resnext50_64x4d": ["layer1.0.conv2.weight", "fc.weight"],
    }
    for name, keys in cases.items():
        m = getattr(models, name)(num_classes=7)
        sd = m.state_dict()
        for k in keys:
            assert k in sd, f"{name}: {k} not in state dict"
        m.train()
        out = m(x)
        assert out.shape == (2, 7), name
        out.sum().backward()
        assert all(p.grad is not None for p in m.parameters()), name
    assert models.resnext50_64x4d().layer1[0].conv2.weight.shape == (256, 4, 3, 3)
    # feature-extractor modes
    assert models.mobilenet_v1(num_classes=0)(x).shape == (2, 1024, 1, 1)
    assert models.mobilenet_v3_large(num_classes=0, with_pool=False, scale=0.5)(x).shape[1] == 480
    assert models.shufflenet_v2_x0_25(num_classes=0, with_pool=False)(x).shape == (2, 512, 2, 2)
    assert models.densenet121(num_classes=0)(x).shape == (2, 1024, 1, 1)
    assert models.squeezenet1_0(num_classes=5)(torch.randn(2, 3, 96, 96)).shape == (2, 5)
    assert models.shufflenet_v2_swish(num_classes=3)(x).shape == (2, 3)
    with pytest.raises(NotImplementedError):
        models.ShuffleNetV2(scale=0.7)
    with pytest.raises(AssertionError):
        models.DenseNet(layers=100)
    with pytest.raises(ValueError):
        models.mobilenet_v1(pretrained=True)


def test_googlenet_and_inception_v3():
    g = models.googlenet(num_classes=11)
    g.eval()
    with torch.no_grad():
        outs = g(torch.randn(1, 3, 224, 224))
    assert isinstance(outs, list) and [tuple(o.shape) for o in outs] == [(1, 11)] * 3
    for k in ("_conv._conv.weight", "_ince3a._conv3r._conv.weight", "_ince5b._convprj._conv.weight", "_fc_out.weight", "_fc_o1.weight", "_out2.bias"):
        assert k in g.state_dict()
    assert g._fc_o1.weight.shape == (1152, 1024)
    inc = models.inception_v3(num_classes=9)
    inc.eval()
    with torch.no_grad():
        assert inc(torch.randn(1, 3, 299, 299)).shape == (1, 9)
    sd = inc.state_dict()
    for k in ("inception_stem.conv_1a_3x3.0.weight", "inception_block_list.0.branch5x5_2.0.weight", "inception_block_list.3.branch3x3.0.weight",
              "inception_block_list.5.branch7x7dbl_5.1.weight", "inception_block_list.8.branch7x7x3_4.0.weight",
              "inception_block_list.10.branch3x3dbl_3b.0.weight", "inception_block_list.10.branch_pool_conv.0.weight", "fc.weight"):
        assert k in sd, k
    assert inc.inception_block_list[4].branch7x7_2[0].weight.shape == (128, 128, 1, 7)
    assert len(inc.inception_block_list) == 11
