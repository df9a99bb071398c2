"""Minimal paddle.vision.transforms (reference: python/paddle/vision/transforms/transforms.py) on tensors/ndarrays."""
import numpy as np
import torch


class Compose:
    def __init__(self, transforms):
        self.transforms = transforms

    def __call__(self, x):
        for t in self.transforms:
            x = t(x)
        return x


class ToTensor:
    def __init__(self, data_format="CHW", keys=None):
        self.data_format = data_format

    def __call__(self, img):
        arr = np.asarray(img)
        if arr.ndim == 2:
            arr = arr[:, :, None]
        t = torch.from_numpy(np.ascontiguousarray(arr)).float()
        if arr.dtype == np.uint8:
            t = t / 255.0
        return t.permute(2, 0, 1) if self.data_format == "CHW" else t


class Normalize:
    def __init__(self, mean=0.0, std=1.0, data_format="CHW", to_rgb=False, keys=None):
        self.mean, self.std, self.data_format = mean, std, data_format

    def __call__(self, img):
        t = img if isinstance(img, torch.Tensor) else torch.as_tensor(np.asarray(img)).float()
        shape = (-1, 1, 1) if self.data_format == "CHW" else (1, 1, -1)
        mean = torch.as_tensor(self.mean, dtype=t.dtype).reshape(shape) if not np.isscalar(self.mean) else self.mean
        std = torch.as_tensor(self.std, dtype=t.dtype).reshape(shape) if not np.isscalar(self.std) else self.std
        return (t - mean) / std


class Resize:
    def __init__(self, size, interpolation="bilinear", keys=None):
        self.size = (size, size) if isinstance(size, int) else tuple(size)

    def __call__(self, img):
        t = img if isinstance(img, torch.Tensor) else ToTensor()(img)
        return torch.nn.functional.interpolate(t[None], size=self.size, mode="bilinear", align_corners=False)[0]


class RandomHorizontalFlip:
    def __init__(self, prob=0.5, keys=None):
        self.prob = prob

    def __call__(self, img):
        if np.random.rand() < self.prob:
            return torch.flip(img, [-1]) if isinstance(img, torch.Tensor) else np.ascontiguousarray(np.asarray(img)[:, ::-1])
        return img


class CenterCrop:
    def __init__(self, size, keys=None):
        self.size = (size, size) if isinstance(size, int) else tuple(size)

    def __call__(self, img):
        t = img if isinstance(img, torch.Tensor) else ToTensor()(img)
        h, w = t.shape[-2:]
        th, tw = self.size
        i, j = (h - th) // 2, (w - tw) // 2
        return t[..., i:i + th, j:j + tw]


class Transpose:
    def __init__(self, order=(2, 0, 1), keys=None):
        self.order = order

    def __call__(self, img):
        return np.asarray(img).transpose(self.order)
