from .resnet import (BasicBlock, BottleneckBlock, ResNet, resnet18, resnet34, resnet50, resnet101, resnet152,  # noqa: F401
                     wide_resnet50_2, resnext50_32x4d)
from .lenet import LeNet  # noqa: F401
from .classic import AlexNet, MobileNetV2, VGG, alexnet, mobilenet_v2, vgg11, vgg13, vgg16, vgg19  # noqa: F401
