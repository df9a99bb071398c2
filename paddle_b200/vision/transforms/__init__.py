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


# ---- further transforms (tensor CHW / HWC ndarray inputs) ------------------------------------------------------
def _chw(img):
    return img if isinstance(img, torch.Tensor) else ToTensor()(img)


class RandomVerticalFlip:
    def __init__(self, prob=0.5, keys=None):
        self.prob = prob

    def __call__(self, img):
        if np.random.rand() < self.prob:
            return torch.flip(img, [-2]) if isinstance(img, torch.Tensor) else np.ascontiguousarray(np.asarray(img)[::-1])
        return img


class Pad:
    def __init__(self, padding, fill=0, padding_mode="constant", keys=None):
        p = [padding] * 4 if isinstance(padding, int) else list(padding)
        self.p = p * 2 if len(p) == 2 else p            # (left, top, right, bottom)
        self.fill, self.mode = fill, padding_mode

    def __call__(self, img):
        t = _chw(img)
        l, tp, r, b = self.p
        mode = {"constant": "constant", "edge": "replicate", "reflect": "reflect", "symmetric": "reflect"}[self.mode]
        kw = dict(value=self.fill) if mode == "constant" else {}
        return torch.nn.functional.pad(t[None], (l, r, tp, b), mode=mode, **kw)[0]


class RandomCrop:
    def __init__(self, size, padding=None, pad_if_needed=False, fill=0, padding_mode="constant", keys=None):
        self.size = (size, size) if isinstance(size, int) else tuple(size)
        self.pad = Pad(padding, fill, padding_mode) if padding is not None else None

    def __call__(self, img):
        t = _chw(img)
        if self.pad is not None:
            t = self.pad(t)
        h, w = t.shape[-2:]
        th, tw = self.size
        i, j = np.random.randint(0, h - th + 1), np.random.randint(0, w - tw + 1)
        return t[..., i:i + th, j:j + tw]


class RandomResizedCrop:
    def __init__(self, size, scale=(0.08, 1.0), ratio=(3.0 / 4, 4.0 / 3), interpolation="bilinear", keys=None):
        self.size = (size, size) if isinstance(size, int) else tuple(size)
        self.scale, self.ratio = scale, ratio

    def __call__(self, img):
        t = _chw(img)
        h, w = t.shape[-2:]
        for _ in range(10):
            area = h * w * np.random.uniform(*self.scale)
            ar = np.exp(np.random.uniform(np.log(self.ratio[0]), np.log(self.ratio[1])))
            cw, ch = int(round(np.sqrt(area * ar))), int(round(np.sqrt(area / ar)))
            if 0 < cw <= w and 0 < ch <= h:
                i, j = np.random.randint(0, h - ch + 1), np.random.randint(0, w - cw + 1)
                t = t[..., i:i + ch, j:j + cw]
                break
        return torch.nn.functional.interpolate(t[None], size=self.size, mode="bilinear", align_corners=False)[0]


class Grayscale:
    def __init__(self, num_output_channels=1, keys=None):
        self.n = num_output_channels

    def __call__(self, img):
        t = _chw(img)
        g = (0.299 * t[0] + 0.587 * t[1] + 0.114 * t[2])[None]
        return g.expand(self.n, -1, -1).contiguous()


class BrightnessTransform:
    def __init__(self, value, keys=None):
        self.value = value

    def __call__(self, img):
        t = _chw(img)
        return (t * np.random.uniform(max(0, 1 - self.value), 1 + self.value)).clamp(0, 1)


class ContrastTransform:
    def __init__(self, value, keys=None):
        self.value = value

    def __call__(self, img):
        t = _chw(img)
        f = np.random.uniform(max(0, 1 - self.value), 1 + self.value)
        m = t.mean()
        return ((t - m) * f + m).clamp(0, 1)


class ColorJitter:
    def __init__(self, brightness=0, contrast=0, saturation=0, hue=0, keys=None):
        self.ts = [BrightnessTransform(brightness)] * bool(brightness) + [ContrastTransform(contrast)] * bool(contrast)

    def __call__(self, img):
        for i in np.random.permutation(len(self.ts)):
            img = self.ts[i](img)
        return img


class RandomErasing:
    def __init__(self, prob=0.5, scale=(0.02, 0.33), ratio=(0.3, 3.3), value=0, inplace=False, keys=None):
        self.prob, self.scale, self.ratio, self.value = prob, scale, ratio, value

    def __call__(self, img):
        t = _chw(img).clone()
        if np.random.rand() >= self.prob:
            return t
        h, w = t.shape[-2:]
        area = h * w * np.random.uniform(*self.scale)
        ar = np.exp(np.random.uniform(np.log(self.ratio[0]), np.log(self.ratio[1])))
        eh, ew = min(h, int(round(np.sqrt(area * ar)))), min(w, int(round(np.sqrt(area / ar))))
        i, j = np.random.randint(0, h - eh + 1), np.random.randint(0, w - ew + 1)
        t[..., i:i + eh, j:j + ew] = self.value
        return t
