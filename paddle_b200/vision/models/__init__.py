"""paddle.vision.models (reference: python/paddle/vision/models/__init__.py)."""
from .classic import AlexNet, MobileNetV2, VGG, alexnet, mobilenet_v2, vgg11, vgg13, vgg16, vgg19  # noqa: F401
from .families import (  # noqa: F401
    DenseNet, GoogLeNet, InceptionV3, MobileNetV1, MobileNetV3Large, MobileNetV3Small, ShuffleNetV2, SqueezeNet, densenet121,
    densenet161, densenet169, densenet201, densenet264, googlenet, inception_v3, mobilenet_v1, mobilenet_v3_large,
    mobilenet_v3_small, shufflenet_v2_swish, shufflenet_v2_x0_25, shufflenet_v2_x0_33, shufflenet_v2_x0_5, shufflenet_v2_x1_0,
    shufflenet_v2_x1_5, shufflenet_v2_x2_0, squeezenet1_0, squeezenet1_1,
)
from .lenet import LeNet  # noqa: F401
from .resnet import (  # noqa: F401
    BasicBlock, BottleneckBlock, ResNet, resnet18, resnet34, resnet50, resnet101, resnet152, resnext50_32x4d, resnext50_64x4d,
    resnext101_32x4d, resnext101_64x4d, resnext152_32x4d, resnext152_64x4d, wide_resnet50_2, wide_resnet101_2,
)

__all__ = [
    "ResNet", "resnet18", "resnet34", "resnet50", "resnet101", "resnet152", "resnext50_32x4d", "resnext50_64x4d", "resnext101_32x4d",
    "resnext101_64x4d", "resnext152_32x4d", "resnext152_64x4d", "wide_resnet50_2", "wide_resnet101_2", "VGG", "vgg11", "vgg13", "vgg16",
    "vgg19", "MobileNetV1", "mobilenet_v1", "MobileNetV2", "mobilenet_v2", "MobileNetV3Small", "MobileNetV3Large", "mobilenet_v3_small",
    "mobilenet_v3_large", "LeNet", "DenseNet", "densenet121", "densenet161", "densenet169", "densenet201", "densenet264", "AlexNet",
    "alexnet", "InceptionV3", "inception_v3", "SqueezeNet", "squeezenet1_0", "squeezenet1_1", "GoogLeNet", "googlenet", "ShuffleNetV2",
    "shufflenet_v2_x0_25", "shufflenet_v2_x0_33", "shufflenet_v2_x0_5", "shufflenet_v2_x1_0", "shufflenet_v2_x1_5", "shufflenet_v2_x2_0",
    "shufflenet_v2_swish",
]
