"""Synthetic stand-ins for paddle.vision.datasets (no network in this environment): same item shapes/dtypes."""
import numpy as np
import torch.utils.data as tud


class _Synthetic(tud.Dataset):
    def __init__(self, shape, num_classes, length, transform=None, seed=0):
        self.shape, self.num_classes, self.length, self.transform = shape, num_classes, length, transform
        self.rng = np.random.RandomState(seed)

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        rng = np.random.RandomState(i)
        img = rng.randint(0, 256, self.shape, dtype=np.uint8)
        label = np.array([rng.randint(0, self.num_classes)], dtype=np.int64)
        if self.transform is not None:
            img = self.transform(img)
        return img, label


class MNIST(_Synthetic):
    def __init__(self, image_path=None, label_path=None, mode="train", transform=None, download=True, backend=None):
        super().__init__((28, 28), 10, 60000 if mode == "train" else 10000, transform)


class Cifar10(_Synthetic):
    def __init__(self, data_file=None, mode="train", transform=None, download=True, backend=None):
        super().__init__((32, 32, 3), 10, 50000 if mode == "train" else 10000, transform)


class FakeImageNet(_Synthetic):
    def __init__(self, length=1281167, transform=None):
        super().__init__((224, 224, 3), 1000, length, transform)
