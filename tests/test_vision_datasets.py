"""paddle.vision.datasets on miniature files in the real formats (reference: python/paddle/vision/datasets/*.py)."""
import gzip
import io
import os
import pickle
import struct
import tarfile

import numpy as np
import pytest
import scipy.io
import torch

import paddle
import paddle.vision.transforms as T
from paddle.vision import datasets as D

Image = pytest.importorskip("PIL.Image")


def _tar_add(tar, name, data):
    info = tarfile.TarInfo(name)
    info.size = len(data)
    tar.addfile(info, io.BytesIO(data))


def _png(arr, mode=None):
    buf = io.BytesIO()
    Image.fromarray(arr, mode).save(buf, format="PNG")
    return buf.getvalue()


def test_mnist_idx_files_and_synthetic_fallback(tmp_path):
    rng = np.random.RandomState(0)
    imgs = rng.randint(0, 256, (7, 28, 28), dtype=np.uint8)
    labels = rng.randint(0, 10, 7).astype(np.uint8)
    ip, lp = str(tmp_path / "train-images-idx3-ubyte.gz"), str(tmp_path / "train-labels-idx1-ubyte.gz")
    with gzip.open(ip, "wb") as f:
        f.write(struct.pack(">HBBIII", 0, 8, 3, 7, 28, 28) + imgs.tobytes())
    with gzip.open(lp, "wb") as f:
        f.write(struct.pack(">HBBI", 0, 8, 1, 7) + labels.tobytes())
    ds = D.MNIST(ip, lp, mode="train")
    assert len(ds) == 7
    img, lab = ds[3]
    assert isinstance(img, Image.Image) and img.mode == "L" and img.size == (28, 28)
    np.testing.assert_array_equal(np.asarray(img), imgs[3])
    assert lab.dtype == np.int64 and lab.tolist() == [int(labels[3])]
    cv = D.FashionMNIST(ip, lp, backend="cv2")
    img, lab = cv[3]
    assert isinstance(img, np.ndarray) and img.dtype == np.float32 and img.shape == (28, 28)
    np.testing.assert_array_equal(img, imgs[3].astype("float32"))
    tens = D.MNIST(ip, lp, transform=T.Compose([T.ToTensor(), T.Normalize([0.5], [0.5])]))[0][0]
    assert isinstance(tens, torch.Tensor) and tens.shape == (1, 28, 28)
    with pytest.raises(ValueError):
        D.MNIST(ip, None)
    with pytest.raises(ValueError):
        D.MNIST(ip, lp, backend="opencv")
    # no files: the synthetic stand-in with the real sizes; deterministic per index
    syn = D.MNIST(mode="test")
    assert len(syn) == 10000 and len(D.MNIST()) == 60000
    a, b = syn[5], D.MNIST(mode="test")[5]
    np.testing.assert_array_equal(np.asarray(a[0]), np.asarray(b[0]))
    assert a[1].tolist() == b[1].tolist() and 0 <= int(a[1][0]) < 10
    loader = paddle.io.DataLoader(D.MNIST(transform=T.ToTensor()), batch_size=4)
    x, y = next(iter(loader))
    assert x.shape == (4, 1, 28, 28) and y.shape == (4, 1)


def test_cifar_tarballs(tmp_path):
    rng = np.random.RandomState(1)
    p10 = str(tmp_path / "cifar-10-python.tar.gz")
    batches = {}
    with tarfile.open(p10, "w:gz") as tar:
        for name, n in (("data_batch_1", 3), ("data_batch_2", 2), ("test_batch", 4)):
            batch = {b"data": rng.randint(0, 256, (n, 3072), dtype=np.uint8), b"labels": rng.randint(0, 10, n).tolist()}
            batches[name] = batch
            _tar_add(tar, f"cifar-10-batches-py/{name}", pickle.dumps(batch))
        _tar_add(tar, "cifar-10-batches-py/batches.meta", pickle.dumps({}))
    train, test = D.Cifar10(p10, "train"), D.Cifar10(p10, "test", backend="cv2")
    assert len(train) == 5 and len(test) == 4
    img, lab = train[3]                                   # first item of data_batch_2
    assert isinstance(img, Image.Image) and img.size == (32, 32) and img.mode == "RGB"
    want = batches["data_batch_2"][b"data"][0].reshape(3, 32, 32).transpose(1, 2, 0)
    np.testing.assert_array_equal(np.asarray(img), want)
    assert lab.tolist() == [batches["data_batch_2"][b"labels"][0]]
    img, lab = test[1]
    assert img.dtype == np.float32 and img.shape == (32, 32, 3)
    p100 = str(tmp_path / "cifar-100-python.tar.gz")
    with tarfile.open(p100, "w:gz") as tar:
        for name, n in (("train", 3), ("test", 2)):
            batch = {b"data": rng.randint(0, 256, (n, 3072), dtype=np.uint8), b"fine_labels": rng.randint(0, 100, n).tolist(),
                     b"coarse_labels": [0] * n}
            batches[name] = batch
            _tar_add(tar, f"cifar-100-python/{name}", pickle.dumps(batch))
    c100 = D.Cifar100(p100, "train")
    assert len(c100) == 3 and c100[2][1].tolist() == [batches["train"][b"fine_labels"][2]] and len(D.Cifar100(p100, "test")) == 2
    assert len(D.Cifar10()) == 50000 and len(D.Cifar100(mode="test")) == 10000 and int(D.Cifar100()[0][1][0]) < 100


def test_folders(tmp_path):
    rng = np.random.RandomState(2)
    for cls, n in (("cat", 2), ("dog", 3)):
        os.makedirs(tmp_path / "root" / cls / "nested", exist_ok=True)
        for i in range(n):
            Image.fromarray(rng.randint(0, 256, (8, 6, 3), dtype=np.uint8)).save(tmp_path / "root" / cls / ("nested" if i else "") / f"{i}.png")
    (tmp_path / "root" / "cat" / "notes.txt").write_text("not an image")
    ds = D.DatasetFolder(str(tmp_path / "root"))
    assert ds.classes == ["cat", "dog"] and ds.class_to_idx == {"cat": 0, "dog": 1}
    assert len(ds) == 5 and ds.targets == [0, 0, 1, 1, 1]
    img, target = ds[4]
    assert isinstance(img, Image.Image) and img.size == (6, 8) and target == 1
    t = D.DatasetFolder(str(tmp_path / "root"), transform=T.Compose([T.Resize((4, 4)), T.ToTensor()]))
    assert t[0][0].shape == (3, 4, 4)
    only0 = D.DatasetFolder(str(tmp_path / "root"), is_valid_file=lambda p: p.endswith("0.png"))
    assert len(only0) == 2
    npy = D.DatasetFolder(str(tmp_path / "root"), loader=lambda p: open(p).read(), extensions=(".txt",))
    assert len(npy) == 1 and npy[0] == ("not an image", 0)
    flat = D.ImageFolder(str(tmp_path / "root"))
    assert len(flat) == 5 and isinstance(flat[0], list) and len(flat[0]) == 1 and flat[0][0].size == (6, 8)
    os.makedirs(tmp_path / "empty" / "a")
    with pytest.raises(RuntimeError):
        D.DatasetFolder(str(tmp_path / "empty"))
    with pytest.raises(RuntimeError):
        D.ImageFolder(str(tmp_path / "empty"))
    # back end switch: cv2 gives BGR arrays
    rgb = np.asarray(ds[0][0])
    paddle.vision.set_image_backend("cv2")
    try:
        assert paddle.vision.get_image_backend() == "cv2"
        arr = D.DatasetFolder(str(tmp_path / "root"))[0][0]
        assert isinstance(arr, np.ndarray)
        np.testing.assert_array_equal(arr, rgb[:, :, ::-1])
        np.testing.assert_array_equal(paddle.vision.image_load(ds.samples[0][0]), rgb[:, :, ::-1])
    finally:
        paddle.vision.set_image_backend("pil")
    assert isinstance(paddle.vision.image_load(ds.samples[0][0]), Image.Image)
    with pytest.raises(ValueError):
        paddle.vision.set_image_backend("gd")


def test_flowers_and_voc(tmp_path):
    rng = np.random.RandomState(3)
    tgz = str(tmp_path / "102flowers.tgz")
    pics = {}
    with tarfile.open(tgz, "w:gz") as tar:
        for i in range(1, 7):
            pics[i] = rng.randint(0, 256, (10, 12, 3), dtype=np.uint8)
            buf = io.BytesIO()
            Image.fromarray(pics[i]).save(buf, format="PNG")       # lossless, so pixels can be compared
            _tar_add(tar, "jpg/image_%05d.jpg" % i, buf.getvalue())
    scipy.io.savemat(str(tmp_path / "imagelabels.mat"), {"labels": np.array([[5, 7, 7, 9, 1, 2]])})
    scipy.io.savemat(str(tmp_path / "setid.mat"), {"trnid": np.array([[1, 2]]), "valid": np.array([[3]]), "tstid": np.array([[4, 5, 6]])})
    args = (tgz, str(tmp_path / "imagelabels.mat"), str(tmp_path / "setid.mat"))
    train, test, valid = D.Flowers(*args, mode="train"), D.Flowers(*args, mode="test"), D.Flowers(*args, mode="valid")
    assert (len(train), len(test), len(valid)) == (3, 2, 1)         # train reads the (larger) tstid list
    img, lab = train[1]
    np.testing.assert_array_equal(np.asarray(img), pics[5])
    assert lab.tolist() == [1] and lab.dtype == np.int64
    img, lab = D.Flowers(*args, mode="test", backend="cv2")[0]
    assert img.dtype == np.float32 and img.shape == (10, 12, 3) and lab.tolist() == [5]
    with pytest.raises(RuntimeError):
        D.Flowers()
    voc = str(tmp_path / "VOCtrainval.tar")
    segs = {}
    with tarfile.open(voc, "w") as tar:
        _tar_add(tar, "VOCdevkit/VOC2012/ImageSets/Segmentation/trainval.txt", b"a\nb\nc\n")
        _tar_add(tar, "VOCdevkit/VOC2012/ImageSets/Segmentation/train.txt", b"a\nb\n")
        _tar_add(tar, "VOCdevkit/VOC2012/ImageSets/Segmentation/val.txt", b"c\n")
        for n in "abc":
            segs[n] = (rng.randint(0, 256, (6, 8, 3), dtype=np.uint8), rng.randint(0, 21, (6, 8), dtype=np.uint8))
            _tar_add(tar, f"VOCdevkit/VOC2012/JPEGImages/{n}.jpg", _png(segs[n][0]))
            _tar_add(tar, f"VOCdevkit/VOC2012/SegmentationClass/{n}.png", _png(segs[n][1]))
    tr, te, va = D.VOC2012(voc, "train"), D.VOC2012(voc, "test"), D.VOC2012(voc, "valid")
    assert (len(tr), len(te), len(va)) == (3, 2, 1)
    img, seg = va[0]
    np.testing.assert_array_equal(np.asarray(img), segs["c"][0])
    np.testing.assert_array_equal(np.asarray(seg), segs["c"][1])
    img, seg = D.VOC2012(voc, "test", backend="cv2")[1]
    assert img.dtype == np.float32 and seg.dtype == np.float32 and seg.shape == (6, 8)
    with pytest.raises(RuntimeError):
        D.VOC2012()
