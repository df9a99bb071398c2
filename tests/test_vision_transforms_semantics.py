"""paddle.vision.transforms against restatements of the documented behaviour (reference:
python/paddle/vision/transforms/{transforms,functional}.py and test/legacy_test/test_transforms.py): the kind of image
that goes in comes out, sizes follow the reference's rules, values follow numpy / PIL / colorsys restatements."""
import colorsys
import random

import numpy as np
import pytest
import torch

import paddle  # noqa: F401
import paddle.vision.transforms as T
import paddle.vision.transforms.functional as F

Image = pytest.importorskip("PIL.Image")
from PIL import ImageEnhance, ImageOps  # noqa: E402


@pytest.fixture
def imgs():
    rng = np.random.RandomState(0)
    a = (rng.rand(10, 14, 3) * 255).astype("uint8")
    return a, Image.fromarray(a), torch.from_numpy(a).permute(2, 0, 1).float() / 255.0


def _as_hwc(x):
    if isinstance(x, torch.Tensor):
        return (x.permute(1, 2, 0) * 255.0).round().numpy()
    return np.asarray(x).astype(np.float64)


def test_kind_is_preserved(imgs):
    a, pil, ten = imgs
    ts = [T.Resize(8), T.Resize((6, 5)), T.CenterCrop(4), T.Pad(2), T.Pad([1, 2], padding_mode="reflect"), T.Grayscale(3),
          T.ColorJitter(0.4, 0.4, 0.4, 0.2), T.RandomResizedCrop(5), T.RandomCrop(6, padding=1), T.RandomRotation(30),
          T.RandomAffine(20, (0.1, 0.1), (0.8, 1.2), 10), T.RandomPerspective(1.0), T.RandomErasing(1.0),
          T.RandomHorizontalFlip(1.0), T.RandomVerticalFlip(1.0), T.BrightnessTransform(0.3), T.HueTransform(0.2)]
    for t in ts:
        oa, op, ot = t(a), t(pil), t(ten)
        assert isinstance(oa, np.ndarray) and oa.dtype == np.uint8 and oa.ndim == 3, t
        assert isinstance(op, Image.Image) and op.mode == "RGB", t
        assert isinstance(ot, torch.Tensor) and ot.dtype == torch.float32 and ot.dim() == 3, t
        # layout: ndarray HWC, tensor CHW, PIL (w, h) - all the same spatial size for deterministic-size transforms
        assert oa.shape[2] == 3 and ot.shape[0] == 3
        assert op.size == (oa.shape[1], oa.shape[0])


def test_resize_rules(imgs):
    a, pil, ten = imgs                                                  # h=10, w=14
    assert T.Resize(5)(a).shape == (5, 7, 3)                            # shorter side -> 5, 14 * 5 / 10 = 7
    assert T.Resize(5)(pil).size == (7, 5)
    assert T.Resize(5)(ten).shape == (3, 5, 7)
    assert F.resize(a.transpose(1, 0, 2).copy(), 5).shape == (7, 5, 3)
    assert T.Resize((3, 9))(a).shape == (3, 9, 3)
    assert F.resize(a, 10) is a                                         # already that size
    # bilinear on arrays: half-pixel centres, no antialias (cv2.INTER_LINEAR's rule)
    g = np.arange(16, dtype=np.float32).reshape(4, 4, 1)
    up = F.resize(g, (8, 8))
    src = np.clip((np.arange(8) + 0.5) / 2 - 0.5, 0, 3)
    i0 = np.floor(src).astype(int)
    i1 = np.minimum(i0 + 1, 3)
    fr = src - i0
    rows = g[i0, :, 0] * (1 - fr)[:, None] + g[i1, :, 0] * fr[:, None]
    want = rows[:, i0] * (1 - fr)[None] + rows[:, i1] * fr[None]
    np.testing.assert_allclose(up[:, :, 0], want, rtol=1e-5, atol=1e-5)
    assert F.resize(g, (2, 2), "nearest").shape == (2, 2, 1)
    with pytest.raises(ValueError):
        F.resize(a, 4, "no-such-mode")


def test_pad_modes_match_numpy(imgs):
    a, pil, ten = imgs
    for mode, npmode in [("constant", "constant"), ("edge", "edge"), ("reflect", "reflect"), ("symmetric", "symmetric")]:
        kw = dict(constant_values=7) if mode == "constant" else {}
        want = np.pad(a, ((2, 4), (1, 3), (0, 0)), mode=npmode, **kw)   # padding = (left 1, top 2, right 3, bottom 4)
        got = F.pad(a, (1, 2, 3, 4), fill=7, padding_mode=mode)
        np.testing.assert_array_equal(got, want)
        np.testing.assert_array_equal(np.asarray(F.pad(pil, (1, 2, 3, 4), fill=7, padding_mode=mode)), want)
        gt = F.pad(ten, (1, 2, 3, 4), fill=7 / 255.0, padding_mode=mode)
        np.testing.assert_allclose(_as_hwc(gt), want, atol=1e-3)
    assert F.pad(a, 3).shape == (16, 20, 3)
    assert F.pad(a, (1, 2)).shape == (14, 16, 3)                        # (left/right, top/bottom)
    np.testing.assert_array_equal(np.asarray(F.pad(pil, 2, fill=9)), np.asarray(ImageOps.expand(pil, 2, (9, 9, 9))))
    got = F.pad(a, 1, fill=(1, 2, 3))
    assert got[0, 0].tolist() == [1, 2, 3]
    with pytest.raises(ValueError):
        F.pad(a, (1, 2, 3))


def test_crop_flip(imgs):
    a, pil, ten = imgs
    np.testing.assert_array_equal(F.crop(a, 2, 3, 4, 5), a[2:6, 3:8])
    np.testing.assert_array_equal(np.asarray(F.crop(pil, 2, 3, 4, 5)), a[2:6, 3:8])
    torch.testing.assert_close(F.crop(ten, 2, 3, 4, 5), ten[:, 2:6, 3:8])
    # centre crop offsets round((H - h) / 2): (10 - 5) / 2 = 2.5 -> 2 (banker's), (14 - 5) / 2 = 4.5 -> 4
    np.testing.assert_array_equal(F.center_crop(a, 5), a[2:7, 4:9])
    np.testing.assert_array_equal(F.center_crop(a, (4, 6)), a[3:7, 4:10])
    np.testing.assert_array_equal(F.hflip(a), a[:, ::-1])
    np.testing.assert_array_equal(F.vflip(a), a[::-1])
    np.testing.assert_array_equal(np.asarray(F.hflip(pil)), a[:, ::-1])
    np.testing.assert_array_equal(np.asarray(F.vflip(pil)), a[::-1])
    torch.testing.assert_close(F.hflip(ten), ten.flip(-1))
    torch.testing.assert_close(F.vflip(ten), ten.flip(-2))
    assert T.RandomHorizontalFlip(0.0)(a) is a
    with pytest.raises(TypeError):
        F.hflip([1, 2, 3])


def test_to_tensor_normalize_transpose(imgs):
    a, pil, ten = imgs
    t = T.ToTensor()(a)
    assert t.shape == (3, 10, 14) and t.dtype == torch.float32
    torch.testing.assert_close(t, ten)
    torch.testing.assert_close(T.ToTensor()(pil), ten)
    assert T.ToTensor("HWC")(a).shape == (10, 14, 3)
    assert T.ToTensor()(a[:, :, 0]).shape == (1, 10, 14)
    f = np.random.rand(4, 5, 3).astype("float32")                       # float arrays are not rescaled
    np.testing.assert_allclose(T.ToTensor()(f).numpy(), f.transpose(2, 0, 1))
    with pytest.raises(ValueError):
        F.to_tensor(a, "WHC")
    mean, std = [0.5, 0.4, 0.3], [0.2, 0.3, 0.4]
    out = T.Normalize(mean, std)(ten)
    want = (ten.numpy() - np.array(mean, "float32")[:, None, None]) / np.array(std, "float32")[:, None, None]
    np.testing.assert_allclose(out.numpy(), want, rtol=1e-6)
    out = T.Normalize(mean=[127.5] * 3, std=[127.5] * 3, data_format="HWC")(a)
    assert isinstance(out, np.ndarray) and out.dtype == np.float32 and out.shape == a.shape
    np.testing.assert_allclose(out, (a.astype("float32") - 127.5) / 127.5, rtol=1e-6)
    out = F.normalize(a, [0.0] * 3, [1.0] * 3, data_format="HWC", to_rgb=True)
    np.testing.assert_array_equal(out, a[..., ::-1].astype("float32"))
    assert isinstance(T.Normalize(0.5, 0.5, data_format="HWC")(pil), np.ndarray)
    tr = T.Transpose()(a)
    assert isinstance(tr, np.ndarray) and tr.shape == (3, 10, 14)
    assert T.Transpose()(pil).shape == (3, 10, 14)
    assert T.Transpose()(a[:, :, 0]).shape == (1, 10, 14)
    c = T.Compose([T.Resize(8), T.CenterCrop(4), T.ToTensor(), T.Normalize([0.5] * 3, [0.5] * 3)])
    assert c(a).shape == (3, 4, 4) and c(pil).shape == (3, 4, 4)
    c = T.Compose([T.RandomResizedCrop(6), T.RandomHorizontalFlip(), T.Transpose(), T.Normalize(127.5, 127.5)])
    o = c(a)
    assert isinstance(o, np.ndarray) and o.shape == (3, 6, 6) and o.dtype == np.float32


def test_color_ops(imgs):
    a, pil, ten = imgs
    af = a.astype(np.float64)
    gray = 0.299 * af[..., 0] + 0.587 * af[..., 1] + 0.114 * af[..., 2]
    # grayscale: same luma as PIL's convert("L") (which truncates a fixed-point sum: +-1)
    g = F.to_grayscale(a)
    assert g.shape == (10, 14, 1) and g.dtype == np.uint8
    assert np.abs(g[..., 0].astype(int) - np.asarray(pil.convert("L")).astype(int)).max() <= 1
    gp = F.to_grayscale(pil)
    assert gp.mode == "L" and np.abs(np.asarray(gp).astype(int) - np.rint(gray)).max() <= 1
    g3 = F.to_grayscale(pil, 3)
    assert g3.mode == "RGB" and (np.asarray(g3)[..., 0] == np.asarray(g3)[..., 2]).all()
    gt = F.to_grayscale(ten, 3)
    assert gt.shape == (3, 10, 14) and torch.equal(gt[0], gt[1])
    with pytest.raises(ValueError):
        F.to_grayscale(a, 2)
    # brightness
    for f in (0.0, 0.6, 1.0, 1.7):
        want = np.clip(af * f, 0, 255)
        assert np.abs(F.adjust_brightness(a, f).astype(float) - want).max() <= 1.0
        assert np.abs(np.asarray(F.adjust_brightness(pil, f)).astype(float)
                      - np.asarray(ImageEnhance.Brightness(pil).enhance(f)).astype(float)).max() <= 1.0
        np.testing.assert_allclose(F.adjust_brightness(ten, f).numpy(), np.clip(ten.numpy() * f, 0, 1), atol=1e-6)
    # contrast: blend with the mean grey level
    for f in (0.0, 0.5, 1.0, 1.6):
        m = np.floor(gray.mean() + 0.5)
        want = np.clip((af - m) * f + m, 0, 255)
        assert np.abs(F.adjust_contrast(a, f).astype(float) - want).max() <= 1.0
        assert np.abs(np.asarray(F.adjust_contrast(pil, f)).astype(float)
                      - np.asarray(ImageEnhance.Contrast(pil).enhance(f)).astype(float)).max() <= 1.0
    mt = (0.299 * ten[0] + 0.587 * ten[1] + 0.114 * ten[2]).mean()
    torch.testing.assert_close(F.adjust_contrast(ten, 0.5), ((ten - mt) * 0.5 + mt).clamp(0, 1))
    # saturation: blend with the grayscale image
    for f in (0.0, 0.5, 1.0, 1.5):
        want = np.clip((af - gray[..., None]) * f + gray[..., None], 0, 255)
        assert np.abs(F.adjust_saturation(a, f).astype(float) - want).max() <= 1.0
        assert np.abs(np.asarray(F.adjust_saturation(pil, f)).astype(float)
                      - np.asarray(ImageEnhance.Color(pil).enhance(f)).astype(float)).max() <= 2.0
    s0 = F.adjust_saturation(ten, 0.0)
    torch.testing.assert_close(s0[0], s0[2])
    for fn in (F.adjust_brightness, F.adjust_contrast, F.adjust_saturation):
        with pytest.raises(ValueError):
            fn(a, -0.1)
    # hue: a rotation of the HSV hue (colorsys restatement)
    np.testing.assert_allclose(F.adjust_hue(a, 0.0).astype(int), a.astype(int), atol=1)
    for hf in (-0.5, -0.2, 0.3, 0.5):
        got = F.adjust_hue(ten, hf).permute(1, 2, 0).numpy()
        src = ten.permute(1, 2, 0).numpy()
        for (y, x) in [(0, 0), (3, 5), (9, 13), (4, 4)]:
            h, s, v = colorsys.rgb_to_hsv(*src[y, x].tolist())
            want = colorsys.hsv_to_rgb((h + hf) % 1.0, s, v)
            np.testing.assert_allclose(got[y, x], want, atol=2e-5)
        assert np.abs(F.adjust_hue(a, hf).astype(float) - got * 255.0).max() <= 1.0
        assert isinstance(F.adjust_hue(pil, hf), Image.Image)
    with pytest.raises(ValueError):
        F.adjust_hue(a, 0.7)
    # ColorJitter uses all four amounts, reproducibly under random.seed
    random.seed(3)
    o1 = T.ColorJitter(0.4, 0.4, 0.4, 0.3)(a)
    random.seed(3)
    o2 = T.ColorJitter(0.4, 0.4, 0.4, 0.3)(a)
    np.testing.assert_array_equal(o1, o2)
    assert not np.array_equal(T.ColorJitter(0, 0, 0.9, 0)(a), a)        # saturation alone changes the image
    assert not np.array_equal(T.ColorJitter(0, 0, 0, 0.4)(a), a)        # hue alone changes the image
    np.testing.assert_array_equal(T.ColorJitter(0, 0, 0, 0)(a), a)
    with pytest.raises(ValueError):
        T.HueTransform(0.7)
    with pytest.raises(ValueError):
        T.BrightnessTransform(-1)


def test_rotate_affine_perspective(imgs):
    a, pil, ten = imgs
    sq = a[:, :10]
    # counter-clockwise quarter turns of a square are exact
    np.testing.assert_array_equal(F.rotate(sq, 90), np.rot90(sq, 1))
    np.testing.assert_array_equal(F.rotate(sq, 180), np.rot90(sq, 2))
    np.testing.assert_array_equal(F.rotate(sq, -90), np.rot90(sq, -1))
    np.testing.assert_array_equal(F.rotate(a, 180), np.rot90(a, 2))
    np.testing.assert_array_equal(F.rotate(a, 0), a)
    # expand: the canvas grows to hold the whole image
    np.testing.assert_array_equal(F.rotate(a, 90, expand=True), np.rot90(a, 1))
    assert F.rotate(pil, 90, expand=True).size == (10, 14)
    np.testing.assert_array_equal(np.asarray(F.rotate(pil, 90, expand=True)), np.rot90(a, 1))
    torch.testing.assert_close(F.rotate(ten, 90, expand=True), torch.rot90(ten, 1, (1, 2)))
    e45 = F.rotate(sq, 45, expand=True)
    assert e45.shape[0] == e45.shape[1] and 14 <= e45.shape[0] <= 16
    # without expand the corners that leave the canvas are filled
    r = F.rotate(a, 90, fill=9)
    assert r.shape == a.shape and (r[:, 0] == 9).all() and (r[:, -1] == 9).all()
    np.testing.assert_array_equal(r[:, 2:12], np.rot90(a, 1)[2:12, :])
    # rotate == PIL's rotate for quarter turns
    np.testing.assert_array_equal(np.asarray(F.rotate(pil, 90, expand=True)), np.asarray(pil.rotate(90, expand=True)))
    # affine: a pure translation moves pixels by (tx, ty) and fills the rest
    out = F.affine(a, 0, (3, 2), 1.0, 0, fill=5)
    np.testing.assert_array_equal(out[2:, 3:], a[:-2, :-3])
    assert (out[:2] == 5).all() and (out[:, :3] == 5).all()
    out = F.affine(a, 0, (-3, 0), 1.0, (0, 0), fill=(1, 2, 3))
    np.testing.assert_array_equal(out[:, :-3], a[:, 3:])
    assert out[0, -1].tolist() == [1, 2, 3]
    # affine's angle is CLOCKWISE: rotate(x) == affine(-x)
    np.testing.assert_array_equal(F.affine(sq, -90, (0, 0), 1.0, 0), np.rot90(sq, 1))
    np.testing.assert_array_equal(F.affine(sq, 30, (0, 0), 1.0, 0), F.rotate(sq, -30))
    # scale 2 about the centre with nearest sampling: the central block, every pixel doubled
    big = F.affine(sq, 0, (0, 0), 2.0, 0)
    assert big.shape == sq.shape
    assert {tuple(p) for p in big.reshape(-1, 3).tolist()} <= {tuple(p) for p in sq[2:8, 2:8].reshape(-1, 3).tolist()}
    # shear along x keeps the centre row in place
    sh = F.affine(sq.astype("float32"), 0, (0, 0), 1.0, (20, 0), "bilinear")
    assert sh.dtype == np.float32 and sh.shape == sq.shape
    with pytest.raises(ValueError):
        F.affine(a, 0, (0, 0), 0.0, 0)
    with pytest.raises(TypeError):
        F.affine(a, 0, 3, 1.0, 0)
    # perspective: identity corners give the image back; a shifted quad is the matching translation
    h, w = 10, 14
    corners = [(0, 0), (w - 1, 0), (w - 1, h - 1), (0, h - 1)]
    np.testing.assert_array_equal(F.perspective(a, corners, corners), a)
    moved = [(x + 2, y + 1) for x, y in corners]
    out = F.perspective(a, corners, moved, fill=4)
    np.testing.assert_array_equal(out[1:, 2:], a[:-1, :-2])
    assert (out[0] == 4).all() and (out[:, :2] == 4).all()
    torch.testing.assert_close(F.perspective(ten, corners, corners, "bilinear"), ten)
    # mapping onto the horizontally mirrored corners is a flip
    mirrored = [(w - 1, 0), (0, 0), (0, h - 1), (w - 1, h - 1)]
    np.testing.assert_array_equal(F.perspective(a, corners, mirrored), a[:, ::-1])


def test_random_transforms_draw_like_the_reference(imgs):
    a, pil, ten = imgs
    random.seed(1)
    np.random.seed(1)
    o1 = T.Compose([T.RandomResizedCrop(6), T.RandomRotation(20), T.RandomCrop(4), T.RandomErasing(1.0)])(a)
    random.seed(1)
    np.random.seed(1)
    o2 = T.Compose([T.RandomResizedCrop(6), T.RandomRotation(20), T.RandomCrop(4), T.RandomErasing(1.0)])(a)
    np.testing.assert_array_equal(o1, o2)
    # RandomCrop: padding, pad_if_needed, size errors
    assert T.RandomCrop(12, pad_if_needed=True)(a).shape == (12, 12, 3)
    assert T.RandomCrop((10, 14))(a) is not None and T.RandomCrop((10, 14))(a).shape == (10, 14, 3)
    with pytest.raises(ValueError):
        T.RandomCrop(20)(a)
    # RandomResizedCrop falls back to a ratio-clamped centre crop when no draw fits
    t = T.RandomResizedCrop(4, scale=(1.0, 1.0), ratio=(0.5, 0.5))
    i, j, h, w = t._get_param(np.zeros((10, 40, 3), "uint8"))          # in_ratio 4 > 0.5 -> h = 10, w = 5
    assert (h, w) == (10, 5) and (i, j) == (0, 17)
    # RandomErasing: value kinds
    e = T.RandomErasing(1.0, value=(1, 2, 3))(a)
    changed = (e != a).any(-1)
    assert changed.any() and {tuple(v) for v in e[changed].tolist()} == {(1, 2, 3)}
    ys, xs = np.where(changed)
    assert changed[ys.min():ys.max() + 1, xs.min():xs.max() + 1].mean() > 0.95   # one rectangle
    et = T.RandomErasing(1.0, value="random")(ten)
    assert et.shape == ten.shape and not torch.equal(et, ten)
    assert T.RandomErasing(0.0)(a) is a
    with pytest.raises(ValueError):
        T.RandomErasing(1.0, value=(1, 2))(a)
    with pytest.raises(ValueError):
        T.RandomErasing(value="noise")
    inp = ten.clone()
    out = T.RandomErasing(1.0, value=0.5, inplace=True)(inp)
    assert out is inp and (inp == 0.5).any()
    # RandomPerspective end points stay inside their corner boxes
    rp = T.RandomPerspective(1.0, 0.5)
    for _ in range(20):
        start, end = rp.get_params(14, 10, 0.5)
        assert start == [(0, 0), (13, 0), (13, 9), (0, 9)]
        (x0, y0), (x1, y1), (x2, y2), (x3, y3) = end
        assert 0 <= x0 <= 3 and 0 <= y0 <= 2 and 10 <= x1 <= 13 and 0 <= y1 <= 2
        assert 10 <= x2 <= 13 and 7 <= y2 <= 9 and 0 <= x3 <= 3 and 7 <= y3 <= 9
    assert T.RandomPerspective(0.0)(a) is a
    # RandomAffine parameter ranges
    ra = T.RandomAffine(15, translate=(0.2, 0.1), scale=(0.5, 1.5), shear=(-5, 5, -2, 2))
    for _ in range(20):
        ang, tr, sc, sh = ra._get_param([14, 10], ra.degrees, ra.translate, ra.scale, ra.shear)
        assert -15 <= ang <= 15 and abs(tr[0]) <= 2.8 and abs(tr[1]) <= 1.0 and 0.5 <= sc <= 1.5
        assert -5 <= sh[0] <= 5 and -2 <= sh[1] <= 2
    with pytest.raises(ValueError):
        T.RandomAffine(-3)
    with pytest.raises(ValueError):
        T.RandomAffine(3, translate=(2, 0))


def test_base_transform_keys():
    """The custom-transform contract from BaseTransform's documentation (transforms.py:125)."""
    class CustomRandomFlip(T.BaseTransform):
        def __init__(self, prob=0.5, keys=None):
            super().__init__(keys)
            self.prob = prob

        def _get_params(self, inputs):
            image = inputs[self.keys.index("image")]
            return {"flip": np.random.random() < self.prob, "size": T._get_image_size(image)}

        def _apply_image(self, image):
            return F.hflip(image) if self.params["flip"] else image

        def _apply_coords(self, coords):
            if self.params["flip"]:
                coords[:, 0] = self.params["size"][0] - coords[:, 0]
            return coords

        def _apply_boxes(self, boxes):
            idxs = np.array([(0, 1), (2, 1), (0, 3), (2, 3)]).flatten()
            coords = np.asarray(boxes).reshape(-1, 4)[:, idxs].reshape(-1, 2)
            coords = self._apply_coords(coords).reshape((-1, 4, 2))
            return np.concatenate((coords.min(axis=1), coords.max(axis=1)), axis=1)

        def _apply_mask(self, mask):
            return F.hflip(mask) if self.params["flip"] else mask

    img = Image.fromarray((np.random.rand(400, 500, 3) * 255.0).astype("uint8"))
    boxes = np.array([[2, 3, 200, 300], [50, 60, 80, 100]])
    mask = img.convert("L")
    out = CustomRandomFlip(1.0)(img)
    np.testing.assert_array_equal(np.asarray(out), np.asarray(img)[:, ::-1])
    oi, ob, om = CustomRandomFlip(1.0, keys=("image", "boxes", "mask"))((img, boxes, mask))
    np.testing.assert_array_equal(ob, [[300, 3, 498, 300], [420, 60, 450, 100]])
    np.testing.assert_array_equal(np.asarray(om), np.asarray(mask)[:, ::-1])
    # elements beyond the keys pass through; an unknown key is rejected at construction
    o = T.Resize(4, keys=("image",))((np.zeros((8, 8, 3), "uint8"), "label"))
    assert o[0].shape == (4, 4, 3) and o[1] == "label"
    with pytest.raises(NotImplementedError):
        T.Resize(4, keys=("image", "nonsense"))
    with pytest.raises(ValueError):
        T.Resize(4, keys=3)
