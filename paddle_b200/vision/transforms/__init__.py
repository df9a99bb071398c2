"""paddle.vision.transforms (reference: python/paddle/vision/transforms/transforms.py, functional.py).

Every transform takes a PIL image, an HWC ``numpy.ndarray`` or a CHW tensor and returns the same kind of image
(``ToTensor`` / ``Normalize`` / ``Transpose`` are the documented exceptions); the math lives in ``functional``.
Random draws come from Python's ``random`` and ``numpy.random`` - the generators the reference's transforms draw from,
so ``random.seed`` / ``np.random.seed`` make a pipeline reproducible the same way.
"""
from __future__ import annotations

import math
import numbers
import random
from collections.abc import Sequence

import numpy as np
import torch

from . import functional
from . import functional as F
from .functional import (  # noqa: F401
    adjust_brightness, adjust_contrast, adjust_hue, adjust_saturation, affine, center_crop, crop, erase, hflip,
    normalize, pad, perspective, resize, rotate, to_grayscale, to_tensor, vflip,
)

__all__ = [
    "BaseTransform", "Compose", "Resize", "RandomResizedCrop", "CenterCrop", "RandomHorizontalFlip",
    "RandomVerticalFlip", "Transpose", "Normalize", "BrightnessTransform", "SaturationTransform", "ContrastTransform",
    "HueTransform", "ColorJitter", "RandomCrop", "Pad", "RandomAffine", "RandomRotation", "RandomPerspective",
    "Grayscale", "ToTensor", "RandomErasing", "to_tensor", "hflip", "vflip", "resize", "pad", "affine", "rotate",
    "perspective", "to_grayscale", "crop", "center_crop", "adjust_brightness", "adjust_contrast", "adjust_hue",
    "normalize", "erase",
]

_get_image_size = F._get_image_size


class Compose:
    """Chain of transforms; a failing stage reports which one failed (transforms.py:72)."""

    def __init__(self, transforms):
        self.transforms = transforms

    def __call__(self, data):
        for f in self.transforms:
            try:
                data = f(data)
            except Exception as e:
                import traceback
                print(f"fail to perform transform [{f}] with error: {e} and stack:\n{traceback.format_exc()}")
                raise
        return data

    def __repr__(self):
        return self.__class__.__name__ + "(" + "".join(f"\n    {t}" for t in self.transforms) + "\n)"


class BaseTransform:
    """``keys`` names the kind of every element of a tuple input ("image", "coords", "boxes", "mask", or a custom
    name with a matching ``_apply_<name>`` method); ``_get_params`` runs once per call and its result is ``self.params``
    while the elements are transformed (transforms.py:125)."""

    def __init__(self, keys=None):
        if keys is None:
            keys = ("image",)
        elif not isinstance(keys, Sequence) or isinstance(keys, str):
            raise ValueError(f"keys should be a sequence, but got keys={keys}")
        for k in keys:
            if self._get_apply(k) is None:
                raise NotImplementedError(f"{k} is unsupported data structure")
        self.keys = keys
        self.params = None

    def _get_params(self, inputs):
        pass

    def __call__(self, inputs):
        if not isinstance(inputs, tuple):
            inputs = (inputs,)
        self.params = self._get_params(inputs)
        outputs = []
        for i in range(min(len(inputs), len(self.keys))):
            apply_func = self._get_apply(self.keys[i])
            outputs.append(inputs[i] if apply_func is None else apply_func(inputs[i]))
        if len(inputs) > len(self.keys):
            outputs.extend(inputs[len(self.keys):])
        return outputs[0] if len(outputs) == 1 else tuple(outputs)

    def _get_apply(self, key):
        return getattr(self, f"_apply_{key}", None)

    def _apply_image(self, image):
        raise NotImplementedError

    def _apply_boxes(self, boxes):
        raise NotImplementedError

    def _apply_mask(self, mask):
        raise NotImplementedError

    def _first_image(self, inputs):
        return inputs[self.keys.index("image")] if "image" in self.keys else inputs[0]


def _size2(size, name="size"):
    if isinstance(size, numbers.Number):
        return (int(size), int(size))
    if isinstance(size, Sequence) and len(size) == 1:
        return (int(size[0]), int(size[0]))
    if not isinstance(size, Sequence) or len(size) != 2:
        raise ValueError(f"{name} should be an int or a sequence of 2 ints, got {size!r}")
    return (int(size[0]), int(size[1]))


class ToTensor(BaseTransform):
    """PIL / ndarray HWC -> float32 tensor in ``data_format`` ("CHW" default); uint8 values are divided by 255."""

    def __init__(self, data_format="CHW", keys=None):
        super().__init__(keys)
        self.data_format = data_format

    def _apply_image(self, img):
        return F.to_tensor(img, self.data_format)


class Resize(BaseTransform):
    """``size`` int: shorter side -> size, aspect kept; (h, w): exactly that."""

    def __init__(self, size, interpolation="bilinear", keys=None):
        super().__init__(keys)
        assert isinstance(size, int) or (isinstance(size, Sequence) and len(size) == 2)
        self.size = size
        self.interpolation = interpolation

    def _apply_image(self, img):
        return F.resize(img, self.size, self.interpolation)


class RandomResizedCrop(BaseTransform):
    """Crop a window of ``scale`` x area and aspect ``ratio`` (10 draws, then a centre crop clamped to the ratio range)
    and resize it to ``size`` (transforms.py:409)."""

    def __init__(self, size, scale=(0.08, 1.0), ratio=(3.0 / 4, 4.0 / 3), interpolation="bilinear", keys=None):
        super().__init__(keys)
        self.size = _size2(size)
        assert scale[0] <= scale[1], "scale should be of kind (min, max)"
        assert ratio[0] <= ratio[1], "ratio should be of kind (min, max)"
        self.scale, self.ratio, self.interpolation = scale, ratio, interpolation

    def _get_param(self, image, attempts=10):
        width, height = _get_image_size(image)
        area = height * width
        for _ in range(attempts):
            target_area = np.random.uniform(*self.scale) * area
            log_ratio = tuple(math.log(x) for x in self.ratio)
            aspect_ratio = math.exp(np.random.uniform(*log_ratio))
            w = int(round(math.sqrt(target_area * aspect_ratio)))
            h = int(round(math.sqrt(target_area / aspect_ratio)))
            if 0 < w <= width and 0 < h <= height:
                return random.randint(0, height - h), random.randint(0, width - w), h, w
        in_ratio = float(width) / float(height)
        if in_ratio < min(self.ratio):
            w = width
            h = int(round(w / min(self.ratio)))
        elif in_ratio > max(self.ratio):
            h = height
            w = int(round(h * max(self.ratio)))
        else:
            w, h = width, height
        return (height - h) // 2, (width - w) // 2, h, w

    def _apply_image(self, img):
        i, j, h, w = self._get_param(img)
        return F.resize(F.crop(img, i, j, h, w), self.size, self.interpolation)


class CenterCrop(BaseTransform):
    def __init__(self, size, keys=None):
        super().__init__(keys)
        self.size = _size2(size)

    def _apply_image(self, img):
        return F.center_crop(img, self.size)


class RandomHorizontalFlip(BaseTransform):
    def __init__(self, prob=0.5, keys=None):
        super().__init__(keys)
        assert 0 <= prob <= 1, "probability must be between 0 and 1"
        self.prob = prob

    def _apply_image(self, img):
        return F.hflip(img) if random.random() < self.prob else img


class RandomVerticalFlip(BaseTransform):
    def __init__(self, prob=0.5, keys=None):
        super().__init__(keys)
        assert 0 <= prob <= 1, "probability must be between 0 and 1"
        self.prob = prob

    def _apply_image(self, img):
        return F.vflip(img) if random.random() < self.prob else img


class Normalize(BaseTransform):
    """``(x - mean) / std``; scalars apply to every channel.  Tensors stay tensors, other images become float32
    ndarrays (transforms.py:774)."""

    def __init__(self, mean=0.0, std=1.0, data_format="CHW", to_rgb=False, keys=None):
        super().__init__(keys)
        if isinstance(mean, numbers.Number):
            mean = [mean, mean, mean]
        if isinstance(std, numbers.Number):
            std = [std, std, std]
        self.mean, self.std, self.data_format, self.to_rgb = mean, std, data_format, to_rgb

    def _apply_image(self, img):
        mean, std = self.mean, self.std
        c = img.shape[0 if self.data_format == "CHW" else -1] if hasattr(img, "shape") and len(img.shape) >= 3 else None
        if c is not None and c != len(mean) and len(set(mean)) == 1 and len(set(std)) == 1:
            mean, std = mean[0], std[0]                                # a scalar given for a non-RGB image
        return F.normalize(img, mean, std, self.data_format, self.to_rgb)


class Transpose(BaseTransform):
    """HWC -> ``order`` (default CHW) as ndarray (tensors are permuted); a 2-D image gets a channel axis first."""

    def __init__(self, order=(2, 0, 1), keys=None):
        super().__init__(keys)
        self.order = order

    def _apply_image(self, img):
        if F._is_tensor_image(img):
            return img.permute(*self.order)
        if F._is_pil_image(img):
            img = np.asarray(img)
        if img.ndim == 2:
            img = img[..., np.newaxis]
        return img.transpose(self.order)


# -------------------------------------------------------------------------------------------------------------- color
def _check_input(value, name, center=1, bound=(0, float("inf")), clip_first_on_zero=True):
    """A jitter amount -> the (low, high) range factors are drawn from (transforms.py:48)."""
    if isinstance(value, numbers.Number):
        if value < 0:
            raise ValueError(f"If {name} is a single number, it must be non negative.")
        value = [center - value, center + value]
        if clip_first_on_zero:
            value[0] = max(value[0], 0)
        if not bound[0] <= value[0] <= value[1] <= bound[1]:
            raise ValueError(f"{name} values should be between {bound}")
    elif isinstance(value, (tuple, list)) and len(value) == 2:
        if not bound[0] <= value[0] <= value[1] <= bound[1]:
            raise ValueError(f"{name} values should be between {bound}")
    else:
        raise TypeError(f"{name} should be a single number or a list/tuple with length 2.")
    if value[0] == value[1] == center:
        value = None
    return value


class BrightnessTransform(BaseTransform):
    def __init__(self, value, keys=None):
        super().__init__(keys)
        self.value = _check_input(value, "brightness")

    def _apply_image(self, img):
        if self.value is None:
            return img
        return F.adjust_brightness(img, random.uniform(self.value[0], self.value[1]))


class ContrastTransform(BaseTransform):
    def __init__(self, value, keys=None):
        super().__init__(keys)
        if isinstance(value, numbers.Number) and value < 0:
            raise ValueError("contrast value should be non-negative")
        self.value = _check_input(value, "contrast")

    def _apply_image(self, img):
        if self.value is None:
            return img
        return F.adjust_contrast(img, random.uniform(self.value[0], self.value[1]))


class SaturationTransform(BaseTransform):
    def __init__(self, value, keys=None):
        super().__init__(keys)
        self.value = _check_input(value, "saturation")

    def _apply_image(self, img):
        if self.value is None:
            return img
        return F.adjust_saturation(img, random.uniform(self.value[0], self.value[1]))


class HueTransform(BaseTransform):
    def __init__(self, value, keys=None):
        super().__init__(keys)
        self.value = _check_input(value, "hue", center=0, bound=(-0.5, 0.5), clip_first_on_zero=False)

    def _apply_image(self, img):
        if self.value is None:
            return img
        return F.adjust_hue(img, random.uniform(self.value[0], self.value[1]))


class ColorJitter(BaseTransform):
    """Brightness, contrast, saturation and hue jitter applied in a random order (transforms.py:1067)."""

    def __init__(self, brightness=0, contrast=0, saturation=0, hue=0, keys=None):
        super().__init__(keys)
        self.brightness, self.contrast, self.saturation, self.hue = brightness, contrast, saturation, hue

    def _get_param(self, brightness, contrast, saturation, hue):
        ts = []
        if brightness is not None:
            ts.append(BrightnessTransform(brightness, self.keys))
        if contrast is not None:
            ts.append(ContrastTransform(contrast, self.keys))
        if saturation is not None:
            ts.append(SaturationTransform(saturation, self.keys))
        if hue is not None:
            ts.append(HueTransform(hue, self.keys))
        random.shuffle(ts)
        return Compose(ts)

    def _apply_image(self, img):
        return self._get_param(self.brightness, self.contrast, self.saturation, self.hue)(img)


class Grayscale(BaseTransform):
    def __init__(self, num_output_channels=1, keys=None):
        super().__init__(keys)
        self.num_output_channels = num_output_channels

    def _apply_image(self, img):
        return F.to_grayscale(img, self.num_output_channels)


# --------------------------------------------------------------------------------------------------------- crop / pad
class Pad(BaseTransform):
    def __init__(self, padding, fill=0, padding_mode="constant", keys=None):
        assert isinstance(padding, (numbers.Number, list, tuple))
        assert isinstance(fill, (numbers.Number, str, list, tuple))
        assert padding_mode in ["constant", "edge", "reflect", "symmetric"]
        if isinstance(padding, Sequence) and len(padding) not in (2, 4):
            raise ValueError(f"Padding must be an int or a 2, or 4 element tuple, not a {len(padding)} element tuple")
        super().__init__(keys)
        self.padding = list(padding) if isinstance(padding, Sequence) else padding
        self.fill, self.padding_mode = fill, padding_mode

    def _apply_image(self, img):
        return F.pad(img, self.padding, self.fill, self.padding_mode)


class RandomCrop(BaseTransform):
    """Random ``size`` window after an optional ``padding``; ``pad_if_needed`` pads an image smaller than the window
    (transforms.py:1155)."""

    def __init__(self, size, padding=None, pad_if_needed=False, fill=0, padding_mode="constant", keys=None):
        super().__init__(keys)
        self.size = _size2(size)
        self.padding, self.pad_if_needed, self.fill, self.padding_mode = padding, pad_if_needed, fill, padding_mode

    def _get_param(self, img, output_size):
        w, h = _get_image_size(img)
        th, tw = output_size
        if h < th or w < tw:
            raise ValueError(f"Required crop size {(th, tw)} is larger than input image size {(h, w)}")
        if w == tw and h == th:
            return 0, 0, h, w
        return random.randint(0, h - th), random.randint(0, w - tw), th, tw

    def _apply_image(self, img):
        if self.padding is not None:
            img = F.pad(img, self.padding, self.fill, self.padding_mode)
        w, h = _get_image_size(img)
        if self.pad_if_needed and w < self.size[1]:
            img = F.pad(img, (self.size[1] - w, 0), self.fill, self.padding_mode)
        if self.pad_if_needed and h < self.size[0]:
            img = F.pad(img, (0, self.size[0] - h), self.fill, self.padding_mode)
        i, j, h, w = self._get_param(img, self.size)
        return F.crop(img, i, j, h, w)


class RandomErasing(BaseTransform):
    """With probability ``prob`` overwrite a random window (``scale`` x area, aspect in ``ratio``) with ``value``: a
    number, one value per channel, or "random" for normal noise (transforms.py:1832)."""

    def __init__(self, prob=0.5, scale=(0.02, 0.33), ratio=(0.3, 3.3), value=0, inplace=False, keys=None):
        super().__init__(keys)
        assert isinstance(scale, (tuple, list)) and isinstance(ratio, (tuple, list))
        assert scale[0] >= 0 and scale[1] <= 1 and scale[0] <= scale[1], "scale should be of kind (min, max) in [0, 1]"
        assert ratio[0] >= 0 and ratio[0] <= ratio[1], "ratio should be of kind (min, max)"
        assert 0 <= prob <= 1, "The probability should be in range [0, 1]"
        assert isinstance(value, (numbers.Number, str, tuple, list))
        if isinstance(value, str) and value != "random":
            raise ValueError("value must be 'random' when type is str")
        self.prob, self.scale, self.ratio, self.value, self.inplace = prob, scale, ratio, value, inplace

    def _get_param(self, img, scale, ratio, value):
        """(top, left, h, w, v); ``v is img`` when no window fitted in 10 draws."""
        if F._is_tensor_image(img):
            c, h, w = img.shape[-3], img.shape[-2], img.shape[-1]
            two_d = False
        else:
            shape = np.asarray(img).shape
            h, w, c = shape[0], shape[1], (shape[2] if len(shape) > 2 else 1)
            two_d = len(shape) == 2
        img_area = h * w
        log_ratio = np.log(ratio)
        for _ in range(10):
            erase_area = np.random.uniform(*scale) * img_area
            aspect_ratio = np.exp(np.random.uniform(*log_ratio))
            erase_h = int(round(np.sqrt(erase_area * aspect_ratio)))
            erase_w = int(round(np.sqrt(erase_area / aspect_ratio)))
            if erase_h >= h or erase_w >= w:
                continue
            if F._is_tensor_image(img):
                if value is None:
                    v = torch.randn(c, erase_h, erase_w).to(img.dtype)
                else:
                    v = torch.as_tensor(value, dtype=img.dtype)[:, None, None]
            else:
                v = np.random.normal(size=[erase_h, erase_w, c]) * 255 if value is None else np.array(value)[None, None, :]
                if two_d:
                    v = v[..., 0]
            top = np.random.randint(0, h - erase_h + 1)
            left = np.random.randint(0, w - erase_w + 1)
            return top, left, erase_h, erase_w, v
        return 0, 0, h, w, img

    def _apply_image(self, img):
        if random.random() >= self.prob:
            return img
        if isinstance(self.value, numbers.Number):
            value = [self.value]
        elif isinstance(self.value, str):
            value = None
        else:
            value = list(self.value)
        c = (img.shape[-3] if F._is_tensor_image(img) else
             (np.asarray(img).shape[2] if np.asarray(img).ndim > 2 else 1))
        if value is not None and len(value) not in (1, c):
            raise ValueError("Value should be a single number or a sequence with length equals to image's channel.")
        if value is not None and len(value) == 1:
            value = value * c
        top, left, eh, ew, v = self._get_param(img, self.scale, self.ratio, value)
        if v is img:
            return img
        return F.erase(img, top, left, eh, ew, v, self.inplace)


# ---------------------------------------------------------------------------------------------------------- geometric
def _check_sequence_input(x, name, req_sizes):
    msg = req_sizes[0] if len(req_sizes) < 2 else " or ".join(str(s) for s in req_sizes)
    if not isinstance(x, Sequence):
        raise TypeError(f"{name} should be a sequence of length {msg}.")
    if len(x) not in req_sizes:
        raise ValueError(f"{name} should be sequence of length {msg}.")


def _setup_angle(x, name, req_sizes=(2,)):
    if isinstance(x, numbers.Number):
        if x < 0:
            raise ValueError(f"If {name} is a single number, it must be positive.")
        x = [-x, x]
    else:
        _check_sequence_input(x, name, req_sizes)
    return [float(d) for d in x]


class RandomAffine(BaseTransform):
    """Random rotation in ``degrees``, translation up to ``translate`` = (fraction of width, fraction of height),
    ``scale`` range, ``shear`` range(s) in degrees (transforms.py:1385)."""

    def __init__(self, degrees, translate=None, scale=None, shear=None, interpolation="nearest", fill=0, center=None,
                 keys=None):
        self.degrees = _setup_angle(degrees, name="degrees", req_sizes=(2,))
        super().__init__(keys)
        assert interpolation in ["nearest", "bilinear", "bicubic"]
        self.interpolation = interpolation
        if translate is not None:
            _check_sequence_input(translate, "translate", req_sizes=(2,))
            for t in translate:
                if not 0.0 <= t <= 1.0:
                    raise ValueError("translation values should be between 0 and 1")
        self.translate = translate
        if scale is not None:
            _check_sequence_input(scale, "scale", req_sizes=(2,))
            for s in scale:
                if s <= 0:
                    raise ValueError("scale values should be positive")
        self.scale = scale
        self.shear = _setup_angle(shear, name="shear", req_sizes=(2, 4)) if shear is not None else None
        if fill is None:
            fill = 0
        elif not isinstance(fill, (Sequence, numbers.Number)):
            raise TypeError("Fill should be either a sequence or a number.")
        self.fill = fill
        if center is not None:
            _check_sequence_input(center, "center", req_sizes=(2,))
        self.center = center

    def _get_param(self, img_size, degrees, translate=None, scale_ranges=None, shears=None):
        angle = random.uniform(degrees[0], degrees[1])
        if translate is not None:
            max_dx = float(translate[0] * img_size[0])
            max_dy = float(translate[1] * img_size[1])
            translations = (int(random.uniform(-max_dx, max_dx)), int(random.uniform(-max_dy, max_dy)))
        else:
            translations = (0, 0)
        scale = random.uniform(scale_ranges[0], scale_ranges[1]) if scale_ranges is not None else 1.0
        shear_x = shear_y = 0.0
        if shears is not None:
            shear_x = random.uniform(shears[0], shears[1])
            if len(shears) == 4:
                shear_y = random.uniform(shears[2], shears[3])
        return angle, translations, scale, (shear_x, shear_y)

    def _apply_image(self, img):
        w, h = _get_image_size(img)
        ret = self._get_param([w, h], self.degrees, self.translate, self.scale, self.shear)
        return F.affine(img, *ret, interpolation=self.interpolation, fill=self.fill, center=self.center)


class RandomRotation(BaseTransform):
    def __init__(self, degrees, interpolation="nearest", expand=False, center=None, fill=0, keys=None):
        if isinstance(degrees, numbers.Number):
            if degrees < 0:
                raise ValueError("If degrees is a single number, it must be positive.")
            self.degrees = (-degrees, degrees)
        else:
            if len(degrees) != 2:
                raise ValueError("If degrees is a sequence, it must be of len 2.")
            self.degrees = tuple(degrees)
        super().__init__(keys)
        self.interpolation, self.expand, self.center, self.fill = interpolation, expand, center, fill

    def _get_param(self, degrees):
        return random.uniform(degrees[0], degrees[1])

    def _apply_image(self, img):
        return F.rotate(img, self._get_param(self.degrees), self.interpolation, self.expand, self.center, self.fill)


class RandomPerspective(BaseTransform):
    """With probability ``prob`` move every corner inwards by up to ``distortion_scale`` x half the side
    (transforms.py:1650)."""

    def __init__(self, prob=0.5, distortion_scale=0.5, interpolation="nearest", fill=0, keys=None):
        super().__init__(keys)
        assert 0 <= prob <= 1, "probability must be between 0 and 1"
        assert 0 <= distortion_scale <= 1, "distortion_scale must be between 0 and 1"
        assert interpolation in ["nearest", "bilinear", "bicubic"]
        assert isinstance(fill, (numbers.Number, str, list, tuple))
        self.prob, self.distortion_scale, self.interpolation, self.fill = prob, distortion_scale, interpolation, fill

    def get_params(self, width, height, distortion_scale):
        """(startpoints, endpoints): the image corners and, for each, a point drawn from the ``distortion_scale`` x
        half-side box inside that corner."""
        dx = int(distortion_scale * (width // 2))
        dy = int(distortion_scale * (height // 2))

        def near(n):                       # [0, n]
            return int(random.uniform(0, n + 1))

        def far(full, n):                  # [full - 1 - n, full - 1]
            return int(random.uniform(full - n - 1, full))
        endpoints = [(near(dx), near(dy)), (far(width, dx), near(dy)), (far(width, dx), far(height, dy)),
                     (near(dx), far(height, dy))]
        startpoints = [(0, 0), (width - 1, 0), (width - 1, height - 1), (0, height - 1)]
        return startpoints, endpoints

    def _apply_image(self, img):
        w, h = _get_image_size(img)
        if random.random() < self.prob:
            start, end = self.get_params(w, h, self.distortion_scale)
            return F.perspective(img, start, end, self.interpolation, self.fill)
        return img
