"""AlexNet / VGG / MobileNetV2 (reference: python/paddle/vision/models/{alexnet,vgg,mobilenetv2}.py), built from this
framework's layers so they train through the same DataParallel / AMP paths as ResNet."""
from __future__ import annotations

from ... import nn


class AlexNet(nn.Layer):
    def __init__(self, num_classes=1000):
        super().__init__()
        self.features = nn.Sequential(
            nn.Conv2D(3, 64, 11, stride=4, padding=2), nn.ReLU(), nn.MaxPool2D(3, 2),
            nn.Conv2D(64, 192, 5, padding=2), nn.ReLU(), nn.MaxPool2D(3, 2),
            nn.Conv2D(192, 384, 3, padding=1), nn.ReLU(), nn.Conv2D(384, 256, 3, padding=1), nn.ReLU(),
            nn.Conv2D(256, 256, 3, padding=1), nn.ReLU(), nn.MaxPool2D(3, 2))
        self.pool = nn.AdaptiveAvgPool2D((6, 6))
        self.classifier = nn.Sequential(nn.Dropout(0.5), nn.Linear(256 * 36, 4096), nn.ReLU(), nn.Dropout(0.5),
                                        nn.Linear(4096, 4096), nn.ReLU(), nn.Linear(4096, num_classes))

    def forward(self, x):
        return self.classifier(self.pool(self.features(x)).flatten(1))


def alexnet(pretrained=False, **kw):
    return AlexNet(**kw)


_VGG = {11: [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
        13: [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
        16: [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
        19: [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"]}


class VGG(nn.Layer):
    def __init__(self, features, num_classes=1000, with_pool=True):
        super().__init__()
        self.features = features
        self.avgpool = nn.AdaptiveAvgPool2D((7, 7)) if with_pool else None
        self.classifier = nn.Sequential(nn.Linear(512 * 49, 4096), nn.ReLU(), nn.Dropout(), nn.Linear(4096, 4096), nn.ReLU(),
                                        nn.Dropout(), nn.Linear(4096, num_classes)) if num_classes > 0 else None

    def forward(self, x):
        x = self.features(x)
        if self.avgpool is not None:
            x = self.avgpool(x)
        return self.classifier(x.flatten(1)) if self.classifier is not None else x


def _vgg(depth, batch_norm=False, **kw):
    layers, c = [], 3
    for v in _VGG[depth]:
        if v == "M":
            layers.append(nn.MaxPool2D(2, 2))
        else:
            layers += [nn.Conv2D(c, v, 3, padding=1)] + ([nn.BatchNorm2D(v)] if batch_norm else []) + [nn.ReLU()]
            c = v
    return VGG(nn.Sequential(*layers), **kw)


def vgg11(pretrained=False, batch_norm=False, **kw):
    return _vgg(11, batch_norm, **kw)


def vgg13(pretrained=False, batch_norm=False, **kw):
    return _vgg(13, batch_norm, **kw)


def vgg16(pretrained=False, batch_norm=False, **kw):
    return _vgg(16, batch_norm, **kw)


def vgg19(pretrained=False, batch_norm=False, **kw):
    return _vgg(19, batch_norm, **kw)


class _InvertedResidual(nn.Layer):
    def __init__(self, inp, oup, stride, expand):
        super().__init__()
        hid = int(round(inp * expand))
        self.use_res = stride == 1 and inp == oup
        layers = []
        if expand != 1:
            layers += [nn.Conv2D(inp, hid, 1, bias_attr=False), nn.BatchNorm2D(hid), nn.ReLU6()]
        layers += [nn.Conv2D(hid, hid, 3, stride=stride, padding=1, groups=hid, bias_attr=False), nn.BatchNorm2D(hid), nn.ReLU6(),
                   nn.Conv2D(hid, oup, 1, bias_attr=False), nn.BatchNorm2D(oup)]
        self.conv = nn.Sequential(*layers)

    def forward(self, x):
        return x + self.conv(x) if self.use_res else self.conv(x)


class MobileNetV2(nn.Layer):
    def __init__(self, scale=1.0, num_classes=1000, with_pool=True):
        super().__init__()
        def md(v, d=8):
            n = max(d, int(v + d / 2) // d * d)
            return n + d if n < 0.9 * v else n
        cfg = [(1, 16, 1, 1), (6, 24, 2, 2), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1)]
        c = md(32 * scale)
        last = md(1280 * max(1.0, scale))
        feats = [nn.Conv2D(3, c, 3, stride=2, padding=1, bias_attr=False), nn.BatchNorm2D(c), nn.ReLU6()]
        for t, ch, n, s in cfg:
            oc = md(ch * scale)
            for i in range(n):
                feats.append(_InvertedResidual(c, oc, s if i == 0 else 1, t))
                c = oc
        feats += [nn.Conv2D(c, last, 1, bias_attr=False), nn.BatchNorm2D(last), nn.ReLU6()]
        self.features = nn.Sequential(*feats)
        self.pool = nn.AdaptiveAvgPool2D(1) if with_pool else None
        self.classifier = nn.Sequential(nn.Dropout(0.2), nn.Linear(last, num_classes)) if num_classes > 0 else None

    def forward(self, x):
        x = self.features(x)
        if self.pool is not None:
            x = self.pool(x)
        return self.classifier(x.flatten(1)) if self.classifier is not None else x


def mobilenet_v2(pretrained=False, scale=1.0, **kw):
    return MobileNetV2(scale=scale, **kw)
