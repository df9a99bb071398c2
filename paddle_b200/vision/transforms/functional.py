"""paddle.vision.transforms.functional (reference: python/paddle/vision/transforms/functional.py and its three back
ends functional_pil.py / functional_cv2.py / functional_tensor.py).

Contract kept from the reference: an image goes in as a PIL image, an HWC (or HW) ``numpy.ndarray`` or a CHW tensor and
comes out as the SAME kind - same layout, same dtype, same value range (uint8 images stay uint8 in [0, 255], float
images stay float).  Only ``to_tensor`` (and ``normalize`` on non-tensors, which returns an ndarray) changes the kind.

One implementation serves the three kinds: the image is viewed as a float32 CHW tensor in its own value range
(``_to_work``), the operation runs on that tensor with torch ops, and ``_from_work`` converts back (round + clip for
integer images).  PIL inputs use PIL's own resampling for ``resize`` so that the result is what a PIL user expects
(antialiased reduction).
"""
from __future__ import annotations

import math
import numbers

import numpy as np
import torch
import torch.nn.functional as TF

try:
    from PIL import Image
except ImportError:                                                    # pragma: no cover
    Image = None

__all__ = []


# ------------------------------------------------------------------------------------------------------ image kinds
def _is_pil_image(img):
    return Image is not None and isinstance(img, Image.Image)


def _is_tensor_image(img):
    return isinstance(img, torch.Tensor)


def _is_numpy_image(img):
    return isinstance(img, np.ndarray) and img.ndim in (2, 3)


def _check(img):
    if not (_is_pil_image(img) or _is_numpy_image(img) or _is_tensor_image(img)):
        raise TypeError(f"img should be PIL Image or Tensor Image or ndarray with dim=[2 or 3]. Got {type(img)}")


def _get_image_size(img):
    """(width, height)."""
    if _is_pil_image(img):
        return img.size
    if _is_numpy_image(img):
        return img.shape[:2][::-1]
    if _is_tensor_image(img):
        return (int(img.shape[-1]), int(img.shape[-2]))
    raise TypeError(f"Unexpected type {type(img)}")


class _Meta:
    __slots__ = ("kind", "dtype", "mode", "gray2d", "vmax", "lead")


def _to_work(img):
    """-> (float32 CHW tensor in the image's own value range, meta)."""
    _check(img)
    m = _Meta()
    m.mode, m.gray2d, m.lead = None, False, None
    if _is_tensor_image(img):
        m.kind, m.dtype = "tensor", img.dtype
        t = img
        if t.dim() == 2:
            m.gray2d, t = True, t[None]
        elif t.dim() > 3:                                              # batch of CHW images: fold the batch into C
            m.lead = tuple(t.shape[:-3])
            t = t.reshape((-1,) + tuple(t.shape[-2:]))
        m.vmax = 1.0 if img.dtype.is_floating_point else 255.0
        return t.to(torch.float32), m
    if _is_pil_image(img):
        m.kind, m.mode = "pil", img.mode
        arr = np.asarray(img)
        if arr.dtype == np.bool_:
            arr = arr.astype(np.uint8) * 255
    else:
        m.kind = "np"
        arr = img
    m.dtype = arr.dtype
    if arr.ndim == 2:
        m.gray2d, arr = True, arr[:, :, None]
    m.vmax = 255.0 if arr.dtype == np.uint8 else (1.0 if arr.dtype.kind == "f" else float(np.iinfo(arr.dtype).max))
    t = torch.from_numpy(np.ascontiguousarray(arr).astype(np.float32, copy=False)).permute(2, 0, 1)
    return t, m


def _from_work(t, m, channels_changed=False):
    if m.kind == "tensor":
        if not m.dtype.is_floating_point:
            t = t.round().clamp(0, m.vmax)
        t = t.to(m.dtype)
        if m.lead is not None:
            per = t.shape[0] // max(int(np.prod(m.lead)), 1)
            t = t.reshape(tuple(m.lead) + (per,) + tuple(t.shape[-2:]))
        elif m.gray2d and t.shape[0] == 1:
            t = t[0]
        return t
    arr = t.permute(1, 2, 0).contiguous().numpy()
    if m.dtype.kind in "ui":
        arr = np.clip(np.rint(arr), 0, m.vmax)
    arr = arr.astype(m.dtype)
    if m.gray2d and arr.shape[2] == 1:
        arr = arr[:, :, 0]
    if m.kind == "np":
        return arr
    if arr.ndim == 3 and arr.shape[2] == 1:
        arr = arr[:, :, 0]
    mode = m.mode
    if channels_changed or mode in ("1", "P"):
        mode = None
    return Image.fromarray(arr, mode) if mode is not None and mode not in ("I", "F") else Image.fromarray(arr)


_PIL_RESAMPLE = {"nearest": 0, "lanczos": 1, "bilinear": 2, "bicubic": 3, "box": 4, "hamming": 5}
_TORCH_MODE = {"nearest": "nearest", "bilinear": "bilinear", "bicubic": "bicubic", "area": "area", "lanczos": "bicubic",
               "box": "area", "hamming": "bilinear"}


# --------------------------------------------------------------------------------------------------------- to_tensor
def to_tensor(pic, data_format="CHW"):
    """PIL / ndarray (HWC) -> float32 tensor, CHW by default; uint8 input is scaled to [0, 1] (functional.py:48)."""
    if data_format not in ("CHW", "HWC"):
        raise ValueError(f"data_format should be CHW or HWC. Got {data_format}")
    if _is_tensor_image(pic):
        return pic
    if not (_is_pil_image(pic) or _is_numpy_image(pic)):
        raise TypeError(f"pic should be PIL Image or Tensor Image or ndarray with dim=[2 or 3]. Got {type(pic)}")
    t, m = _to_work(pic)
    if m.kind == "pil" and m.mode in ("I", "I;16", "F"):
        pass                                                           # integer / float modes keep their values
    elif m.dtype == np.uint8:
        t = t / 255.0
    return t.contiguous() if data_format == "CHW" else t.permute(1, 2, 0).contiguous()


# ------------------------------------------------------------------------------------------------------------ resize
def _resize_target(w, h, size):
    if isinstance(size, numbers.Integral):
        size = int(size)
        if (w <= h and w == size) or (h <= w and h == size):
            return h, w
        if w < h:
            return int(size * h / w), size
        return size, int(size * w / h)
    if not (isinstance(size, (list, tuple)) and len(size) == 2):
        raise TypeError(f"Got inappropriate size arg: {size}")
    return int(size[0]), int(size[1])


def resize(img, size, interpolation="bilinear"):
    """``size`` int: the SHORTER side becomes ``size`` and the aspect ratio is kept; (h, w): exact (functional.py:100)."""
    _check(img)
    w, h = _get_image_size(img)
    oh, ow = _resize_target(w, h, size)
    if _is_pil_image(img):
        if interpolation not in _PIL_RESAMPLE:
            raise ValueError(f"unsupported interpolation {interpolation!r} for PIL images")
        return img.resize((ow, oh), _PIL_RESAMPLE[interpolation])
    if interpolation not in _TORCH_MODE:
        raise ValueError(f"unsupported interpolation {interpolation!r}")
    if (oh, ow) == (h, w):
        return img
    t, m = _to_work(img)
    mode = _TORCH_MODE[interpolation]
    kw = dict(align_corners=False) if mode in ("bilinear", "bicubic") else {}
    out = TF.interpolate(t[None], size=(oh, ow), mode=mode, **kw)[0]
    if mode == "bicubic":
        out = out.clamp(0, m.vmax) if m.kind != "tensor" or not m.dtype.is_floating_point else out
    return _from_work(out, m)


# --------------------------------------------------------------------------------------------------- pad, crop, flip
def _pad4(padding):
    if isinstance(padding, numbers.Integral):
        return [int(padding)] * 4
    if not isinstance(padding, (list, tuple)) or len(padding) not in (2, 4):
        raise ValueError(f"Padding must be an int or a 2, or 4 element tuple, not {padding!r}")
    p = [int(v) for v in padding]
    return [p[0], p[1], p[0], p[1]] if len(p) == 2 else p              # (left, top, right, bottom)


def pad(img, padding, fill=0, padding_mode="constant"):
    """``padding``: int (all sides), (left/right, top/bottom) or (left, top, right, bottom) (functional.py:156).
    ``edge`` repeats the border, ``reflect`` mirrors without repeating it, ``symmetric`` mirrors repeating it."""
    _check(img)
    if padding_mode not in ("constant", "edge", "reflect", "symmetric"):
        raise ValueError("Padding mode should be either constant, edge, reflect or symmetric")
    l, tp, r, b = _pad4(padding)
    t, m = _to_work(img)
    c, h, w = t.shape
    if padding_mode == "constant":
        if isinstance(fill, (list, tuple)):
            if len(fill) != c:
                raise ValueError(f"fill has {len(fill)} values for an image with {c} channels")
            out = torch.empty((c, h + tp + b, w + l + r), dtype=t.dtype)
            out[:] = torch.tensor(fill, dtype=t.dtype).reshape(c, 1, 1)
            out[:, tp:tp + h, l:l + w] = t
        else:
            out = TF.pad(t, (l, r, tp, b), mode="constant", value=float(fill))
    else:
        # index maps, so that a pad wider than the image (legal for symmetric) still works
        def index(n, before, after, mode):
            i = np.arange(-before, n + after)
            if mode == "edge":
                return np.clip(i, 0, n - 1)
            if mode == "reflect":
                if n == 1:
                    return np.zeros_like(i)
                period = 2 * (n - 1)
                i = np.mod(i, period)
                return np.where(i >= n, period - i, i)
            period = 2 * n                                             # symmetric
            i = np.mod(i, period)
            return np.where(i >= n, period - 1 - i, i)
        iy = torch.from_numpy(index(h, tp, b, padding_mode)).long()
        ix = torch.from_numpy(index(w, l, r, padding_mode)).long()
        out = t[:, iy][:, :, ix]
    return _from_work(out, m)


def crop(img, top, left, height, width):
    """The window [top, top + height) x [left, left + width) (functional.py:220)."""
    _check(img)
    if _is_pil_image(img):
        return img.crop((left, top, left + width, top + height))
    if _is_tensor_image(img):
        return img[..., top:top + height, left:left + width]
    return img[top:top + height, left:left + width]


def center_crop(img, output_size):
    _check(img)
    if isinstance(output_size, numbers.Number):
        output_size = (int(output_size), int(output_size))
    w, h = _get_image_size(img)
    th, tw = output_size
    top = int(round((h - th) / 2.0))
    left = int(round((w - tw) / 2.0))
    return crop(img, top, left, th, tw)


def hflip(img):
    _check(img)
    if _is_pil_image(img):
        return img.transpose(Image.FLIP_LEFT_RIGHT)
    if _is_tensor_image(img):
        return torch.flip(img, [-1])
    return np.ascontiguousarray(img[:, ::-1])


def vflip(img):
    _check(img)
    if _is_pil_image(img):
        return img.transpose(Image.FLIP_TOP_BOTTOM)
    if _is_tensor_image(img):
        return torch.flip(img, [-2])
    return np.ascontiguousarray(img[::-1])


# -------------------------------------------------------------------------------------------------------------- color
def _gray(t):
    """Luma of a CHW work tensor (ITU-R 601-2, what PIL's ``convert('L')`` and cv2's RGB2GRAY use)."""
    if t.shape[0] == 1:
        return t
    if t.shape[0] < 3:
        raise ValueError("grayscale conversion needs 1 or 3 channels")
    return (0.299 * t[0] + 0.587 * t[1] + 0.114 * t[2])[None]


def _blend(t, other, factor, m):
    return (factor * t + (1.0 - factor) * other).clamp(0, m.vmax)


def to_grayscale(img, num_output_channels=1):
    """1 or 3 (r == g == b) channel luma image of the same kind (functional.py:888)."""
    if num_output_channels not in (1, 3):
        raise ValueError("num_output_channels should be either 1 or 3")
    t, m = _to_work(img)
    g = _gray(t)
    if num_output_channels == 3:
        g = g.expand(3, -1, -1)
    if m.kind == "np" and num_output_channels == 1:
        m.gray2d = False                                               # ndarray result keeps a channel axis: (H, W, 1)
    return _from_work(g.contiguous(), m, channels_changed=(g.shape[0] != t.shape[0]))


def adjust_brightness(img, brightness_factor):
    """0 -> black, 1 -> the image, 2 -> twice as bright; clipped to the value range (functional.py:374)."""
    if brightness_factor < 0:
        raise ValueError("brightness_factor is not non-negative.")
    t, m = _to_work(img)
    return _from_work(_blend(t, torch.zeros_like(t), brightness_factor, m), m)


def adjust_contrast(img, contrast_factor):
    """Blend with the mean grey level: 0 -> solid grey, 1 -> the image (functional.py:426)."""
    if contrast_factor < 0:
        raise ValueError("contrast_factor is not non-negative.")
    t, m = _to_work(img)
    mean = _gray(t).mean()
    if m.kind != "tensor" and m.dtype.kind in "ui":
        mean = torch.floor(mean + 0.5)                                 # the integer back ends round the mean
    return _from_work(_blend(t, mean.expand_as(t), contrast_factor, m), m)


def adjust_saturation(img, saturation_factor):
    """Blend with the grayscale image: 0 -> black and white, 1 -> the image (functional.py:465)."""
    if saturation_factor < 0:
        raise ValueError("saturation_factor is not non-negative.")
    t, m = _to_work(img)
    if t.shape[0] == 1:
        return img
    return _from_work(_blend(t, _gray(t).expand_as(t), saturation_factor, m), m)


def _rgb_to_hsv(t):
    r, g, b = t[0], t[1], t[2]
    mx, mn = t.max(0).values, t.min(0).values
    d = mx - mn
    nz = d > 0
    dd = torch.where(nz, d, torch.ones_like(d))
    rc, gc, bc = [(mx - c) / dd for c in (r, g, b)]
    h = torch.zeros_like(mx)
    h = torch.where((mx == b) & nz, 4.0 + gc - rc, h)
    h = torch.where((mx == g) & nz, 2.0 + rc - bc, h)
    h = torch.where((mx == r) & nz, bc - gc, h)
    h = (h / 6.0) % 1.0
    s = torch.where(mx > 0, d / torch.where(mx > 0, mx, torch.ones_like(mx)), torch.zeros_like(mx))
    return torch.stack([h, s, mx])


def _hsv_to_rgb(t):
    h, s, v = t[0], t[1], t[2]
    i = torch.floor(h * 6.0)
    f = h * 6.0 - i
    p, q, u = v * (1 - s), v * (1 - f * s), v * (1 - (1 - f) * s)
    i = i.long() % 6

    def sel(*a):
        out = a[5]
        for k in range(4, -1, -1):
            out = torch.where(i == k, a[k], out)
        return out
    return torch.stack([sel(v, q, p, p, u, v), sel(u, v, v, q, p, p), sel(p, p, u, v, v, q)])


def adjust_hue(img, hue_factor):
    """Shift the hue channel by ``hue_factor`` of a full turn, -0.5 <= hue_factor <= 0.5 (functional.py:505)."""
    if not -0.5 <= hue_factor <= 0.5:
        raise ValueError(f"hue_factor:{hue_factor} is not in [-0.5, 0.5].")
    t, m = _to_work(img)
    if t.shape[0] == 1:
        return img
    hsv = _rgb_to_hsv(t[:3] / m.vmax)
    hsv[0] = (hsv[0] + hue_factor) % 1.0
    rgb = _hsv_to_rgb(hsv) * m.vmax
    if t.shape[0] > 3:
        rgb = torch.cat([rgb, t[3:]], 0)
    return _from_work(rgb.clamp(0, m.vmax), m)


# ------------------------------------------------------------------------------------------------- normalize, erase
def normalize(img, mean, std, data_format="CHW", to_rgb=False):
    """``(img - mean) / std`` per channel.  Tensors stay tensors; PIL images and ndarrays give a float32 ndarray in
    ``data_format`` layout semantics (the array is taken to be in that layout already) (functional.py:929)."""
    if _is_tensor_image(img):
        shape = (-1, 1, 1) if data_format == "CHW" else (1, 1, -1)
        dt = img.dtype if img.dtype.is_floating_point else torch.float32
        mean_t = torch.as_tensor(mean, dtype=dt, device=img.device)
        std_t = torch.as_tensor(std, dtype=dt, device=img.device)
        if mean_t.dim() > 0:
            mean_t = mean_t.reshape(shape)
        if std_t.dim() > 0:
            std_t = std_t.reshape(shape)
        return (img.to(dt) - mean_t) / std_t
    arr = np.asarray(img).astype(np.float32)
    if to_rgb:
        arr = arr[..., ::-1] if data_format == "HWC" else arr[::-1]
    shape = (-1, 1, 1) if data_format == "CHW" else (1, 1, -1)
    mean_a = np.asarray(mean, dtype=np.float32)
    std_a = np.asarray(std, dtype=np.float32)
    if mean_a.ndim > 0:
        mean_a = mean_a.reshape(shape)
    if std_a.ndim > 0:
        std_a = std_a.reshape(shape)
    return (arr - mean_a) / std_a


def erase(img, i, j, h, w, v, inplace=False):
    """Write ``v`` into the window [i, i + h) x [j, j + w) (functional.py:969)."""
    _check(img)
    if _is_tensor_image(img):
        out = img if inplace else img.clone()
        out[..., i:i + h, j:j + w] = torch.as_tensor(v, dtype=out.dtype, device=out.device) if not np.isscalar(v) else v
        return out
    if _is_pil_image(img):
        arr = np.array(img)
        arr[i:i + h, j:j + w] = np.asarray(v).astype(arr.dtype) if not np.isscalar(v) else v
        return Image.fromarray(arr, img.mode)
    out = img if inplace else img.copy()
    out[i:i + h, j:j + w] = v
    return out


# ---------------------------------------------------------------------------------------------------------- geometric
def _sample(t, src_x, src_y, interpolation, fill, m):
    """Output pixel (y, x) takes the input at (src_y[y, x], src_x[y, x]) (pixel-index coordinates); points outside the
    input get ``fill``."""
    if interpolation not in ("nearest", "bilinear", "bicubic"):
        raise ValueError(f"unsupported interpolation {interpolation!r}")
    c, h, w = t.shape
    gx = 2.0 * src_x / max(w - 1, 1) - 1.0 if w > 1 else torch.zeros_like(src_x)
    gy = 2.0 * src_y / max(h - 1, 1) - 1.0 if h > 1 else torch.zeros_like(src_y)
    grid = torch.stack([gx, gy], -1)[None].to(torch.float32)
    src = torch.cat([t, torch.ones((1, h, w), dtype=t.dtype)], 0)[None]
    out = TF.grid_sample(src, grid, mode=interpolation, padding_mode="zeros", align_corners=True)[0]
    img, mask = out[:c], out[c:].clamp(0, 1)
    if interpolation == "nearest":
        # grid_sample's nearest keeps points up to half a pixel outside; the reference's back ends drop them
        inside = ((src_x > -0.5) & (src_x < w - 0.5) & (src_y > -0.5) & (src_y < h - 0.5)).to(t.dtype)[None]
        img, mask = img * inside, mask * inside
    if isinstance(fill, (list, tuple)):
        fv = torch.tensor([float(f) for f in fill], dtype=t.dtype)
        fv = (fv if fv.numel() == c else fv[:1].expand(c)).reshape(c, 1, 1)
    else:
        fv = float(fill)
    res = img + (1.0 - mask) * fv
    if interpolation == "bicubic":
        res = res.clamp(0, m.vmax) if not (m.kind == "tensor" and m.dtype.is_floating_point) else res
    return res


def _affine_matrix(center, angle, translate, scale, shear):
    """Forward 3x3 matrix in pixel coordinates: T(translate) C R(angle, clockwise) Sh(shear) S(scale) C^-1
    (reference: functional.py:554 _get_affine_matrix)."""
    a = math.radians(angle)
    sx, sy = math.radians(shear[0]), math.radians(shear[1])
    cx, cy = center
    m00 = math.cos(a - sy) / math.cos(sy)
    m01 = -math.cos(a - sy) * math.tan(sx) / math.cos(sy) - math.sin(a)
    m10 = math.sin(a - sy) / math.cos(sy)
    m11 = -math.sin(a - sy) * math.tan(sx) / math.cos(sy) + math.cos(a)
    rss = np.array([[m00 * scale, m01 * scale, 0.0], [m10 * scale, m11 * scale, 0.0], [0.0, 0.0, 1.0]])
    to_c = np.array([[1.0, 0.0, cx], [0.0, 1.0, cy], [0.0, 0.0, 1.0]])
    from_c = np.array([[1.0, 0.0, -cx], [0.0, 1.0, -cy], [0.0, 0.0, 1.0]])
    tr = np.array([[1.0, 0.0, translate[0]], [0.0, 1.0, translate[1]], [0.0, 0.0, 1.0]])
    return tr @ to_c @ rss @ from_c


def _apply_matrix(img, fwd, out_hw, interpolation, fill):
    t, m = _to_work(img)
    inv = np.linalg.inv(fwd)
    oh, ow = out_hw
    ys, xs = torch.meshgrid(torch.arange(oh, dtype=torch.float64), torch.arange(ow, dtype=torch.float64), indexing="ij")
    src_x = inv[0, 0] * xs + inv[0, 1] * ys + inv[0, 2]
    src_y = inv[1, 0] * xs + inv[1, 1] * ys + inv[1, 2]
    # snap values a rounding error away from a pixel centre, so that exact rotations stay exact
    src_x = torch.where((src_x - src_x.round()).abs() < 1e-6, src_x.round(), src_x).to(torch.float32)
    src_y = torch.where((src_y - src_y.round()).abs() < 1e-6, src_y.round(), src_y).to(torch.float32)
    return _from_work(_sample(t, src_x, src_y, interpolation, fill, m), m)


def affine(img, angle, translate, scale, shear, interpolation="nearest", fill=0, center=None):
    """Rotation by ``angle`` degrees CLOCKWISE about ``center`` (default: the image centre), then ``translate`` =
    (horizontal, vertical) pixels, isotropic ``scale`` and ``shear`` = (x, y) degrees (functional.py:585)."""
    _check(img)
    if not isinstance(angle, (int, float)):
        raise TypeError("Argument angle should be int or float")
    if not isinstance(translate, (list, tuple)) or len(translate) != 2:
        raise TypeError("Argument translate should be a sequence of length 2")
    if scale <= 0.0:
        raise ValueError("Argument scale should be positive")
    if isinstance(shear, numbers.Number):
        shear = [float(shear), 0.0]
    shear = [float(s) for s in shear]
    if len(shear) == 1:
        shear = [shear[0], shear[0]]
    if len(shear) != 2:
        raise ValueError(f"Shear should be a sequence containing two values. Got {shear}")
    w, h = _get_image_size(img)
    c = ((w - 1) * 0.5, (h - 1) * 0.5) if center is None else (float(center[0]), float(center[1]))
    fwd = _affine_matrix(c, float(angle), (float(translate[0]), float(translate[1])), float(scale), shear)
    return _apply_matrix(img, fwd, (h, w), interpolation, fill)


def rotate(img, angle, interpolation="nearest", expand=False, center=None, fill=0):
    """Rotation by ``angle`` degrees COUNTER-clockwise; ``expand`` grows the canvas to hold the whole rotated image
    (functional.py:719)."""
    _check(img)
    w, h = _get_image_size(img)
    c = ((w - 1) * 0.5, (h - 1) * 0.5) if center is None else (float(center[0]), float(center[1]))
    fwd = _affine_matrix(c, -float(angle), (0.0, 0.0), 1.0, (0.0, 0.0))
    oh, ow = h, w
    if expand:
        # corners of the pixel grid (edges at -0.5 / n - 0.5) through the forward map
        pts = np.array([[-0.5, -0.5, 1.0], [w - 0.5, -0.5, 1.0], [w - 0.5, h - 0.5, 1.0], [-0.5, h - 0.5, 1.0]]).T
        q = fwd @ pts
        x0, x1 = q[0].min(), q[0].max()
        y0, y1 = q[1].min(), q[1].max()
        ow = int(math.ceil(x1 - x0 - 1e-6))
        oh = int(math.ceil(y1 - y0 - 1e-6))
        shift = np.array([[1.0, 0.0, (ow - 1) * 0.5 - (x0 + x1) * 0.5], [0.0, 1.0, (oh - 1) * 0.5 - (y0 + y1) * 0.5],
                          [0.0, 0.0, 1.0]])
        fwd = shift @ fwd
    return _apply_matrix(img, fwd, (oh, ow), interpolation, fill)


def _perspective_coeffs(startpoints, endpoints):
    """The 8 coefficients of the map output pixel -> input pixel that sends every end point to its start point
    (reference: functional.py:785)."""
    a, b = [], []
    for (sx, sy), (ex, ey) in zip(startpoints, endpoints):
        a += [[ex, ey, 1, 0, 0, 0, -sx * ex, -sx * ey], [0, 0, 0, ex, ey, 1, -sy * ex, -sy * ey]]
        b += [sx, sy]
    return np.linalg.lstsq(np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64), rcond=None)[0]


def perspective(img, startpoints, endpoints, interpolation="nearest", fill=0):
    """Projective warp taking the four ``startpoints`` (x, y) to the four ``endpoints`` (functional.py:830)."""
    _check(img)
    t, m = _to_work(img)
    h, w = t.shape[-2:]
    k = _perspective_coeffs(startpoints, endpoints)
    ys, xs = torch.meshgrid(torch.arange(h, dtype=torch.float64), torch.arange(w, dtype=torch.float64), indexing="ij")
    den = k[6] * xs + k[7] * ys + 1.0
    src_x = ((k[0] * xs + k[1] * ys + k[2]) / den)
    src_y = ((k[3] * xs + k[4] * ys + k[5]) / den)
    src_x = torch.where((src_x - src_x.round()).abs() < 1e-6, src_x.round(), src_x).to(torch.float32)
    src_y = torch.where((src_y - src_y.round()).abs() < 1e-6, src_y.round(), src_y).to(torch.float32)
    return _from_work(_sample(t, src_x, src_y, interpolation, fill, m), m)
