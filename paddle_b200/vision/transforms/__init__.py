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


# ---- functional API + remaining classes (reference: vision/transforms/functional.py, transforms.py) ---------------
class BaseTransform:
    def __init__(self, keys=None):
        self.keys = keys

    def _apply_image(self, img):
        raise NotImplementedError

    def __call__(self, inputs):
        return self._apply_image(inputs)


def to_tensor(pic, data_format="CHW"):
    return ToTensor(data_format)(pic)


def hflip(img):
    return torch.flip(_chw(img), [-1])


def vflip(img):
    return torch.flip(_chw(img), [-2])


def resize(img, size, interpolation="bilinear"):
    return Resize(size, interpolation)(img)


def pad(img, padding, fill=0, padding_mode="constant"):
    return Pad(padding, fill, padding_mode)(img)


def crop(img, top, left, height, width):
    return _chw(img)[..., top:top + height, left:left + width]


def center_crop(img, output_size):
    return CenterCrop(output_size)(img)


def to_grayscale(img, num_output_channels=1):
    return Grayscale(num_output_channels)(img)


def normalize(img, mean, std, data_format="CHW", to_rgb=False):
    return Normalize(mean, std, data_format)(img)


def erase(img, i, j, h, w, v, inplace=False):
    t = _chw(img) if inplace else _chw(img).clone()
    t[..., i:i + h, j:j + w] = v
    return t


def adjust_brightness(img, brightness_factor):
    return (_chw(img) * brightness_factor).clamp(0, 1)


def adjust_contrast(img, contrast_factor):
    t = _chw(img)
    m = to_grayscale(t).mean() if t.shape[0] == 3 else t.mean()
    return ((t - m) * contrast_factor + m).clamp(0, 1)


def _rgb_to_hsv(t):
    r, g, b = t[0], t[1], t[2]
    mx, mn = t.max(0).values, t.min(0).values
    d = mx - mn
    h = torch.zeros_like(mx)
    nz = d > 0
    rc, gc, bc = [(mx - c) / torch.where(nz, d, torch.ones_like(d)) for c in (r, g, b)]
    h = torch.where((mx == r) & nz, bc - gc, h)
    h = torch.where((mx == g) & nz, 2.0 + rc - bc, h)
    h = torch.where((mx == b) & nz & (mx != r) & (mx != g), 4.0 + gc - rc, h)
    h = (h / 6.0) % 1.0
    s = torch.where(mx > 0, d / torch.where(mx > 0, mx, torch.ones_like(mx)), torch.zeros_like(mx))
    return torch.stack([h, s, mx])


def _hsv_to_rgb(t):
    h, s, v = t[0], t[1], t[2]
    i = torch.floor(h * 6.0)
    f = h * 6.0 - i
    p, q, u = v * (1 - s), v * (1 - f * s), v * (1 - (1 - f) * s)
    i = i.long() % 6
    sel = lambda a0, a1, a2, a3, a4, a5: torch.where(i == 0, a0, torch.where(i == 1, a1, torch.where(i == 2, a2, torch.where(  # noqa: E731
        i == 3, a3, torch.where(i == 4, a4, a5)))))
    return torch.stack([sel(v, q, p, p, u, v), sel(u, v, v, q, p, p), sel(p, p, u, v, v, q)])


def adjust_hue(img, hue_factor):
    hsv = _rgb_to_hsv(_chw(img))
    hsv[0] = (hsv[0] + hue_factor) % 1.0
    return _hsv_to_rgb(hsv)


def adjust_saturation(img, saturation_factor):
    t = _chw(img)
    g = to_grayscale(t)
    return ((t - g) * saturation_factor + g).clamp(0, 1)


def affine(img, angle, translate, scale, shear, interpolation="nearest", fill=0, center=None):
    import math
    t = _chw(img)
    h, w = t.shape[-2:]
    a = math.radians(angle)
    sx, sy = (math.radians(s) for s in (shear if isinstance(shear, (list, tuple)) else (shear, 0.0)))
    # forward matrix (output = M * input), inverted for grid sampling; coordinates normalised to [-1, 1]
    m = torch.tensor([[math.cos(a - sy) / math.cos(sy), -math.cos(a - sy) * math.tan(sx) / math.cos(sy) - math.sin(a), 0.0],
                      [math.sin(a - sy) / math.cos(sy), -math.sin(a - sy) * math.tan(sx) / math.cos(sy) + math.cos(a), 0.0],
                      [0.0, 0.0, 1.0]]) * scale
    m[2, 2] = 1.0
    m[0, 2], m[1, 2] = 2.0 * translate[0] / w, 2.0 * translate[1] / h
    inv = torch.linalg.inv(m)[:2]
    grid = torch.nn.functional.affine_grid(inv[None], [1, t.shape[0], h, w], align_corners=False)
    mode = "bilinear" if interpolation == "bilinear" else "nearest"
    return torch.nn.functional.grid_sample(t[None].float(), grid, mode=mode, padding_mode="zeros", align_corners=False)[0].to(t.dtype)


def rotate(img, angle, interpolation="nearest", expand=False, center=None, fill=0):
    return affine(img, -angle, (0, 0), 1.0, (0.0, 0.0), interpolation, fill, center)


def perspective(img, startpoints, endpoints, interpolation="nearest", fill=0):
    t = _chw(img)
    h, w = t.shape[-2:]
    a, bvec = [], []
    for (sx, sy), (ex, ey) in zip(startpoints, endpoints):
        a += [[ex, ey, 1, 0, 0, 0, -sx * ex, -sx * ey], [0, 0, 0, ex, ey, 1, -sy * ex, -sy * ey]]
        bvec += [sx, sy]
    coef = torch.linalg.lstsq(torch.tensor(a, dtype=torch.float64), torch.tensor(bvec, dtype=torch.float64)[:, None]).solution[:, 0].float()
    ys, xs = torch.meshgrid(torch.arange(h, dtype=torch.float32), torch.arange(w, dtype=torch.float32), indexing="ij")
    den = coef[6] * xs + coef[7] * ys + 1
    gx = (coef[0] * xs + coef[1] * ys + coef[2]) / den
    gy = (coef[3] * xs + coef[4] * ys + coef[5]) / den
    grid = torch.stack([2 * gx / max(w - 1, 1) - 1, 2 * gy / max(h - 1, 1) - 1], -1)[None]
    mode = "bilinear" if interpolation == "bilinear" else "nearest"
    return torch.nn.functional.grid_sample(t[None].float(), grid, mode=mode, padding_mode="zeros", align_corners=True)[0].to(t.dtype)


class SaturationTransform:
    def __init__(self, value, keys=None):
        self.value = value

    def __call__(self, img):
        return adjust_saturation(img, np.random.uniform(max(0, 1 - self.value), 1 + self.value))


class HueTransform:
    def __init__(self, value, keys=None):
        self.value = value

    def __call__(self, img):
        return adjust_hue(img, np.random.uniform(-self.value, self.value))


class RandomRotation:
    def __init__(self, degrees, interpolation="nearest", expand=False, center=None, fill=0, keys=None):
        self.degrees = (-degrees, degrees) if isinstance(degrees, (int, float)) else tuple(degrees)
        self.interpolation = interpolation

    def __call__(self, img):
        return rotate(img, np.random.uniform(*self.degrees), self.interpolation)


class RandomAffine:
    def __init__(self, degrees, translate=None, scale=None, shear=None, interpolation="nearest", fill=0, center=None, keys=None):
        self.degrees = (-degrees, degrees) if isinstance(degrees, (int, float)) else tuple(degrees)
        self.translate, self.scale, self.shear, self.interpolation = translate, scale, shear, interpolation

    def __call__(self, img):
        t = _chw(img)
        h, w = t.shape[-2:]
        tr = (0, 0) if self.translate is None else (np.random.uniform(-self.translate[0], self.translate[0]) * w,
                                                     np.random.uniform(-self.translate[1], self.translate[1]) * h)
        sc = 1.0 if self.scale is None else np.random.uniform(*self.scale)
        sh = (0.0, 0.0) if self.shear is None else (np.random.uniform(-self.shear, self.shear) if isinstance(self.shear, (int, float))
                                                    else np.random.uniform(self.shear[0], self.shear[1]), 0.0)
        return affine(t, np.random.uniform(*self.degrees), tr, sc, sh, self.interpolation)


class RandomPerspective:
    def __init__(self, prob=0.5, distortion_scale=0.5, interpolation="nearest", fill=0, keys=None):
        self.prob, self.d, self.interpolation = prob, distortion_scale, interpolation

    def __call__(self, img):
        t = _chw(img)
        if np.random.rand() >= self.prob:
            return t
        h, w = t.shape[-2:]
        dx, dy = int(self.d * w / 2), int(self.d * h / 2)
        r = lambda n: np.random.randint(0, n + 1)  # noqa: E731
        start = [(0, 0), (w - 1, 0), (w - 1, h - 1), (0, h - 1)]
        end = [(r(dx), r(dy)), (w - 1 - r(dx), r(dy)), (w - 1 - r(dx), h - 1 - r(dy)), (r(dx), h - 1 - r(dy))]
        return perspective(t, start, end, self.interpolation)
