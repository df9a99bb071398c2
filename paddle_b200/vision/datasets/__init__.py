"""paddle.vision.datasets (reference: python/paddle/vision/datasets/{folder,mnist,cifar,flowers,voc2012}.py).

Every dataset parses the same files the reference downloads - MNIST's idx archives, the CIFAR python tarballs, the
Oxford flowers tarball + .mat files, the VOC2012 tarball, class-per-directory image folders - and returns items in the
reference's layout: ``backend="pil"`` gives PIL images, ``"cv2"`` HWC ndarrays (read through PIL here; cv2 is not
installed), labels are int64 arrays of shape [1].

There is no network in this environment.  When no file is given, MNIST / FashionMNIST / Cifar10 / Cifar100 fall back
to a deterministic synthetic stand-in with the real shapes, dtypes, lengths and class counts (so training scripts and
benchmarks run); ``FakeImageNet`` is the ImageNet-shaped stand-in the benchmarks use.  The other datasets need their
files and say so.
"""
from __future__ import annotations

import gzip
import io
import os
import pickle
import struct
import tarfile

import numpy as np
import torch

from ...io import Dataset

try:
    from PIL import Image
except ImportError:                                                    # pragma: no cover
    Image = None

__all__ = ["DatasetFolder", "ImageFolder", "MNIST", "FashionMNIST", "Flowers", "Cifar10", "Cifar100", "VOC2012"]

IMG_EXTENSIONS = (".jpg", ".jpeg", ".png", ".ppm", ".bmp", ".pgm", ".tif", ".tiff", ".webp")

_image_backend = ["pil"]


def set_image_backend(backend):
    if backend not in ("pil", "cv2", "tensor"):
        raise ValueError(f"Expected backend are one of ['pil', 'cv2', 'tensor'], but got {backend}")
    _image_backend[0] = backend


def get_image_backend():
    return _image_backend[0]


def image_load(path, backend=None):
    """PIL image (``pil``) or HWC ndarray (``cv2``: BGR channel order, as OpenCV would give)."""
    backend = backend or _image_backend[0]
    if backend not in ("pil", "cv2", "tensor"):
        raise ValueError(f"Expected backend are one of ['pil', 'cv2', 'tensor'], but got {backend}")
    img = Image.open(path)
    if backend == "pil":
        return img
    arr = np.asarray(img.convert("RGB"))
    if backend == "cv2":
        return np.ascontiguousarray(arr[:, :, ::-1])
    return torch.from_numpy(arr.copy())


def _backend(backend):
    backend = backend or _image_backend[0]
    if backend not in ("pil", "cv2"):
        raise ValueError(f"Expected backend are one of ['pil', 'cv2'], but got {backend}")
    return backend


def _default_dtype():
    return {torch.float32: "float32", torch.float64: "float64", torch.float16: "float16"}.get(torch.get_default_dtype(), "float32")


# ------------------------------------------------------------------------------------------------------------ folders
def has_valid_extension(filename, extensions):
    assert isinstance(extensions, (list, tuple)), "`extensions` must be list or tuple."
    return filename.lower().endswith(tuple(x.lower() for x in extensions))


def make_dataset(dir, class_to_idx, extensions, is_valid_file=None):
    dir = os.path.expanduser(dir)
    if extensions is not None:
        def is_valid_file(x):  # noqa: F811
            return has_valid_extension(x, extensions)
    images = []
    for target in sorted(class_to_idx):
        d = os.path.join(dir, target)
        if not os.path.isdir(d):
            continue
        for root, _, fnames in sorted(os.walk(d, followlinks=True)):
            for fname in sorted(fnames):
                path = os.path.join(root, fname)
                if is_valid_file(path):
                    images.append((path, class_to_idx[target]))
    return images


def pil_loader(path):
    with open(path, "rb") as f:
        return Image.open(f).convert("RGB")


def cv2_loader(path):
    return np.ascontiguousarray(np.asarray(pil_loader(path))[:, :, ::-1])


def default_loader(path):
    return cv2_loader(path) if _image_backend[0] == "cv2" else pil_loader(path)


class DatasetFolder(Dataset):
    """``root/<class>/**/<file>``: classes are the sorted sub-directory names, items are (loaded sample, class index)."""

    def __init__(self, root, loader=None, extensions=None, transform=None, is_valid_file=None):
        self.root, self.transform = root, transform
        if extensions is None and is_valid_file is None:
            extensions = IMG_EXTENSIONS
        classes, class_to_idx = self._find_classes(root)
        samples = make_dataset(root, class_to_idx, extensions, is_valid_file)
        if not samples:
            raise RuntimeError(f"Found 0 directories in subfolders of: {root}\nSupported extensions are: "
                               f"{','.join(extensions or ())}")
        self.loader = default_loader if loader is None else loader
        self.extensions, self.classes, self.class_to_idx = extensions, classes, class_to_idx
        self.samples = samples
        self.targets = [s[1] for s in samples]
        self.dtype = _default_dtype()

    @staticmethod
    def _find_classes(dir):
        classes = sorted(d.name for d in os.scandir(dir) if d.is_dir())
        return classes, {c: i for i, c in enumerate(classes)}

    def __getitem__(self, index):
        path, target = self.samples[index]
        sample = self.loader(path)
        if self.transform is not None:
            sample = self.transform(sample)
        return sample, target

    def __len__(self):
        return len(self.samples)


class ImageFolder(Dataset):
    """Every image under ``root`` (recursively, sorted), no labels: items are ``[sample]``."""

    def __init__(self, root, loader=None, extensions=None, transform=None, is_valid_file=None):
        self.root = root
        if extensions is None and is_valid_file is None:
            extensions = IMG_EXTENSIONS
        if extensions is not None:
            def is_valid_file(x):  # noqa: F811
                return has_valid_extension(x, extensions)
        samples = []
        for base, _, fnames in sorted(os.walk(os.path.expanduser(root), followlinks=True)):
            for fname in sorted(fnames):
                f = os.path.join(base, fname)
                if is_valid_file(f):
                    samples.append(f)
        if not samples:
            raise RuntimeError(f"Found 0 files in subfolders of: {root}\nSupported extensions are: {','.join(extensions or ())}")
        self.loader = default_loader if loader is None else loader
        self.extensions, self.samples, self.transform = extensions, samples, transform

    def __getitem__(self, index):
        sample = self.loader(self.samples[index])
        if self.transform is not None:
            sample = self.transform(sample)
        return [sample]

    def __len__(self):
        return len(self.samples)


# --------------------------------------------------------------------------------------------------- synthetic stand-in
class _Synthetic:
    """Deterministic images of a given shape: item i is generated from seed i."""

    def __init__(self, shape, num_classes, length):
        self.shape, self.num_classes, self.length = shape, num_classes, length

    def image(self, i):
        return np.random.RandomState(i).randint(0, 256, self.shape, dtype=np.uint8)

    def label(self, i):
        return int(np.random.RandomState(i + 7919).randint(0, self.num_classes))


def _finish(image, label, transform, backend, dtype, mode=None):
    """Common tail of ``__getitem__``: PIL for ``pil``, transform, dtype cast for ``cv2``."""
    if backend == "pil":
        image = Image.fromarray(image, mode=mode) if mode else Image.fromarray(image)
    if transform is not None:
        image = transform(image)
    label = np.array([label]).astype("int64")
    if backend == "pil":
        return image, label
    return (image.astype(dtype) if isinstance(image, np.ndarray) else image), label


# --------------------------------------------------------------------------------------------------------------- MNIST
def _read_idx(path):
    """An idx file (optionally gzipped): magic 0x00 0x00 <dtype> <ndim>, big-endian dims, then the data."""
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rb") as f:
        zero, dtype_code, ndim = struct.unpack(">HBB", f.read(4))
        if zero != 0 or dtype_code != 0x08:
            raise ValueError(f"{path} is not an unsigned-byte idx file")
        dims = struct.unpack(">" + "I" * ndim, f.read(4 * ndim))
        data = np.frombuffer(f.read(), dtype=np.uint8)
    return data.reshape(dims)


class MNIST(Dataset):
    """``image_path`` / ``label_path``: the idx3 / idx1 (``.gz``) files.  Items: (28 x 28 image, int64[1])."""
    NAME = "mnist"
    _LENGTHS = {"train": 60000, "test": 10000}

    def __init__(self, image_path=None, label_path=None, mode="train", transform=None, download=True, backend=None):
        assert mode.lower() in ("train", "test"), f"mode should be 'train' or 'test', but got {mode}"
        self.mode, self.transform, self.backend = mode.lower(), transform, _backend(backend)
        self.dtype = _default_dtype()
        self.image_path, self.label_path = image_path, label_path
        if image_path is None and label_path is None:
            self._synthetic = _Synthetic((28, 28), 10, self._LENGTHS[self.mode])
            self.images = self.labels = None
        else:
            if image_path is None or label_path is None:
                raise ValueError("image_path and label_path must be given together")
            self._synthetic = None
            self.images, self.labels = _read_idx(image_path), _read_idx(label_path)
            if self.images.ndim != 3 or len(self.images) != len(self.labels):
                raise ValueError("image / label files do not match")

    def __getitem__(self, idx):
        if self._synthetic is not None:
            image, label = self._synthetic.image(idx), self._synthetic.label(idx)
        else:
            image, label = self.images[idx], int(self.labels[idx])
        return _finish(np.ascontiguousarray(image), label, self.transform, self.backend, self.dtype, "L")

    def __len__(self):
        return self._synthetic.length if self._synthetic is not None else len(self.labels)


class FashionMNIST(MNIST):
    NAME = "fashion-mnist"


# --------------------------------------------------------------------------------------------------------------- CIFAR
class Cifar10(Dataset):
    """``data_file``: ``cifar-10-python.tar.gz``.  Batches whose member name contains ``data_batch`` are the training
    split, ``test_batch`` the test split.  Items: (32 x 32 x 3 image, int64[1])."""
    _TRAIN, _TEST, _LABEL_KEY, _CLASSES = "data_batch", "test_batch", b"labels", 10
    _LENGTHS = {"train": 50000, "test": 10000}

    def __init__(self, data_file=None, mode="train", transform=None, download=True, backend=None):
        assert mode.lower() in ("train", "test"), f"mode.lower() should be 'train' or 'test', but got {mode}"
        self.mode, self.transform, self.backend = mode.lower(), transform, _backend(backend)
        self.dtype = _default_dtype()
        self.data_file = data_file
        if data_file is None:
            self._synthetic = _Synthetic((32, 32, 3), self._CLASSES, self._LENGTHS[self.mode])
            self.data = None
        else:
            self._synthetic = None
            self._load_data()

    def _load_data(self):
        flag = self._TRAIN if self.mode == "train" else self._TEST
        self.data = []
        with tarfile.open(self.data_file, mode="r") as f:
            for name in sorted(n for n in f.getnames() if flag in n):
                batch = pickle.load(f.extractfile(name), encoding="bytes")
                labels = batch.get(self._LABEL_KEY, batch.get(b"labels", batch.get(b"fine_labels")))
                assert labels is not None, f"no labels in {name}"
                self.data += list(zip(batch[b"data"], labels))

    def __getitem__(self, idx):
        if self._synthetic is not None:
            image, label = self._synthetic.image(idx), self._synthetic.label(idx)
        else:
            flat, label = self.data[idx]
            image = np.ascontiguousarray(np.asarray(flat, dtype=np.uint8).reshape(3, 32, 32).transpose(1, 2, 0))
        return _finish(image, int(label), self.transform, self.backend, self.dtype)

    def __len__(self):
        return self._synthetic.length if self._synthetic is not None else len(self.data)


class Cifar100(Cifar10):
    """``cifar-100-python.tar.gz``: members ``train`` / ``test``, fine labels."""
    _TRAIN, _TEST, _LABEL_KEY, _CLASSES = "train", "test", b"fine_labels", 100


class FakeImageNet(Dataset):
    """ImageNet-shaped synthetic data for benchmarks: (224 x 224 x 3 uint8, int64[1]) over 1000 classes."""

    def __init__(self, length=1281167, transform=None):
        self._synthetic, self.transform = _Synthetic((224, 224, 3), 1000, length), transform

    def __getitem__(self, idx):
        image = self._synthetic.image(idx)
        if self.transform is not None:
            image = self.transform(image)
        return image, np.array([self._synthetic.label(idx)], dtype=np.int64)

    def __len__(self):
        return self._synthetic.length


# ------------------------------------------------------------------------------------------------------------- Flowers
class Flowers(Dataset):
    """Oxford 102 flowers: ``data_file`` = 102flowers.tgz (``jpg/image_%05d.jpg``), ``label_file`` = imagelabels.mat,
    ``setid_file`` = setid.mat.  As in the reference the larger ``tstid`` list is the TRAINING split and ``trnid``
    the test split.  Images are read straight from the tarball (no extraction next to it)."""
    MODE_FLAG_MAP = {"train": "tstid", "test": "trnid", "valid": "valid"}

    def __init__(self, data_file=None, label_file=None, setid_file=None, mode="train", transform=None, download=True,
                 backend=None):
        assert mode.lower() in ("train", "valid", "test"), f"mode should be 'train', 'valid' or 'test', but got {mode}"
        self.backend, self.transform = _backend(backend), transform
        for name, value in (("data_file", data_file), ("label_file", label_file), ("setid_file", setid_file)):
            if not value:
                raise RuntimeError(f"Flowers: {name} is not set and this environment has no network access to download it")
        import scipy.io
        self.labels = scipy.io.loadmat(label_file)["labels"][0]
        self.indexes = scipy.io.loadmat(setid_file)[self.MODE_FLAG_MAP[mode.lower()]][0]
        self._tar = tarfile.open(data_file)
        self._members = {m.name.lstrip("./"): m for m in self._tar.getmembers()}
        self.dtype = _default_dtype()

    def __getitem__(self, idx):
        index = int(self.indexes[idx])
        raw = self._tar.extractfile(self._members["jpg/image_%05d.jpg" % index]).read()
        image = Image.open(io.BytesIO(raw))
        image.load()
        if self.backend == "cv2":
            image = np.array(image)
        if self.transform is not None:
            image = self.transform(image)
        label = np.array([self.labels[index - 1]]).astype("int64")
        if self.backend == "pil":
            return image, label
        return (image.astype(self.dtype) if isinstance(image, np.ndarray) else image), label

    def __len__(self):
        return len(self.indexes)

    def __del__(self):
        tar = getattr(self, "_tar", None)
        if tar is not None:
            tar.close()


# ----------------------------------------------------------------------------------------------------------------- VOC
class VOC2012(Dataset):
    """Segmentation pairs from ``VOCtrainval_11-May-2012.tar``: (JPEG image, class-index PNG).  As in the reference,
    ``train`` reads the ``trainval`` list, ``test`` the ``train`` list and ``valid`` the ``val`` list."""
    SET_FILE = "VOCdevkit/VOC2012/ImageSets/Segmentation/{}.txt"
    DATA_FILE = "VOCdevkit/VOC2012/JPEGImages/{}.jpg"
    LABEL_FILE = "VOCdevkit/VOC2012/SegmentationClass/{}.png"
    MODE_FLAG_MAP = {"train": "trainval", "test": "train", "valid": "val"}

    def __init__(self, data_file=None, mode="train", transform=None, download=True, backend=None):
        assert mode.lower() in ("train", "valid", "test"), f"mode should be 'train', 'valid' or 'test', but got {mode}"
        self.backend, self.transform = _backend(backend), transform
        if data_file is None:
            raise RuntimeError("VOC2012: data_file is not set and this environment has no network access to download it")
        self.flag, self.data_file = self.MODE_FLAG_MAP[mode.lower()], data_file
        self.dtype = _default_dtype()
        self.data_tar = tarfile.open(data_file)
        self.name2mem = {m.name: m for m in self.data_tar.getmembers()}
        names = [line.strip().decode("utf-8") for line in self.data_tar.extractfile(self.name2mem[self.SET_FILE.format(self.flag)])]
        self.data = [self.DATA_FILE.format(n) for n in names if n]
        self.labels = [self.LABEL_FILE.format(n) for n in names if n]

    def __getitem__(self, idx):
        data = Image.open(io.BytesIO(self.data_tar.extractfile(self.name2mem[self.data[idx]]).read()))
        label = Image.open(io.BytesIO(self.data_tar.extractfile(self.name2mem[self.labels[idx]]).read()))
        if self.backend == "cv2":
            data, label = np.array(data), np.array(label)
        if self.transform is not None:
            data = self.transform(data)
        if self.backend == "cv2":
            return data.astype(self.dtype), label.astype(self.dtype)
        return data, label

    def __len__(self):
        return len(self.data)

    def __del__(self):
        tar = getattr(self, "data_tar", None)
        if tar is not None:
            tar.close()
