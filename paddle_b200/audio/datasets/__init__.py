"""paddle.audio.datasets (reference: python/paddle/audio/datasets/{dataset,esc50,tess}.py).

``AudioClassificationDataset`` turns (file, label) lists into (feature, label) items; ``ESC50`` and ``TESS`` index the
extracted archives.  There is no network here, so the archives are never downloaded: they are looked up under
``data_dir`` or ``~/.cache/paddle/dataset/audio`` and a clear error says where to put them."""
from __future__ import annotations

import collections
import os

import numpy as np

from ...io import Dataset
from .. import backends
from ..features import MFCC, LogMelSpectrogram, MelSpectrogram, Spectrogram

__all__ = ["ESC50", "TESS"]

feat_funcs = {"raw": None, "melspectrogram": MelSpectrogram, "mfcc": MFCC, "logmelspectrogram": LogMelSpectrogram,
              "spectrogram": Spectrogram}

DATA_HOME = os.path.expanduser(os.path.join("~", ".cache", "paddle", "dataset", "audio"))


class AudioClassificationDataset(Dataset):
    def __init__(self, files, labels, feat_type="raw", sample_rate=None, **kwargs):
        super().__init__()
        if feat_type not in feat_funcs:
            raise RuntimeError(f"Unknown feat_type: {feat_type}, it must be one in {list(feat_funcs.keys())}")
        self.files, self.labels, self.feat_type, self.sample_rate, self.feat_config = files, labels, feat_type, sample_rate, kwargs
        self._extractor = None

    def _convert_to_record(self, idx):
        file, label = self.files[idx], self.labels[idx]
        waveform, sample_rate = backends.load(file)
        self.sample_rate = sample_rate
        feat_func = feat_funcs[self.feat_type]
        record = {}
        if feat_func is not None:
            if self._extractor is None:
                self._extractor = feat_func(sr=sample_rate, **self.feat_config) if self.feat_type != "spectrogram" \
                    else feat_func(**self.feat_config)
            record["feat"] = self._extractor(waveform).squeeze(0)
        else:
            record["feat"] = waveform.squeeze(0)
        record["label"] = label
        return record

    def __getitem__(self, idx):
        record = self._convert_to_record(idx)
        return record["feat"], np.array(record["label"], dtype=np.int64)

    def __len__(self):
        return len(self.files)


def _require(path, what):
    if not os.path.exists(path):
        raise RuntimeError(f"{what} was not found at {path}: this environment has no network access, extract the archive there "
                           f"(or pass data_dir=...)")
    return path


class ESC50(AudioClassificationDataset):
    """2000 five-second environmental recordings, 50 classes, 5 folds; fold ``split`` is the dev set."""
    audio_path = os.path.join("ESC-50-master", "audio")
    meta = os.path.join("ESC-50-master", "meta", "esc50.csv")
    meta_info = collections.namedtuple("META_INFO", ("filename", "fold", "target", "category", "esc10", "src_file", "take"))

    def __init__(self, mode="train", split=1, feat_type="raw", archive=None, data_dir=None, **kwargs):
        assert split in range(1, 6), f"The selected split should be integer, and 1 <= split <= 5, but got {split}"
        self.root = data_dir or DATA_HOME
        files, labels = self._get_data(mode, split)
        super().__init__(files=files, labels=labels, feat_type=feat_type, **kwargs)

    def _get_meta_info(self):
        with open(_require(os.path.join(self.root, self.meta), "ESC-50 meta file")) as rf:
            return [self.meta_info(*line.strip().split(",")) for line in rf.readlines()[1:]]

    def _get_data(self, mode, split):
        files, labels = [], []
        for sample in self._get_meta_info():
            in_dev = int(sample.fold) == split
            if (mode == "train") != in_dev:
                files.append(os.path.join(self.root, self.audio_path, sample.filename))
                labels.append(int(sample.target))
        return files, labels


class TESS(AudioClassificationDataset):
    """2800 recordings of 200 words in 7 emotions by 2 actresses; ``n_folds`` round-robin folds."""
    audio_path = "TESS_Toronto_emotional_speech_set"
    label_list = ["angry", "disgust", "fear", "happy", "neutral", "ps", "sad"]
    meta_info = collections.namedtuple("META_INFO", ("speaker", "word", "emotion"))

    def __init__(self, mode="train", n_folds=5, split=1, feat_type="raw", archive=None, data_dir=None, **kwargs):
        assert isinstance(n_folds, int) and n_folds >= 1, f"the n_folds should be integer and n_folds >= 1, but got {n_folds}"
        assert split in range(1, n_folds + 1), f"The selected split should be integer and should be 1 <= split <= {n_folds}, but got {split}"
        self.root = data_dir or DATA_HOME
        files, labels = self._get_data(mode, n_folds, split)
        super().__init__(files=files, labels=labels, feat_type=feat_type, **kwargs)

    def _get_data(self, mode, n_folds, split):
        wav_files = []
        for root, _, names in sorted(os.walk(_require(os.path.join(self.root, self.audio_path), "TESS audio directory"))):
            wav_files += [os.path.join(root, n) for n in sorted(names) if n.endswith(".wav")]
        files, labels = [], []
        for idx, path in enumerate(wav_files):
            emotion = self.meta_info(*os.path.basename(path)[:-4].split("_")).emotion
            fold = idx % n_folds + 1
            if (mode == "train") != (fold == split):
                files.append(path)
                labels.append(self.label_list.index(emotion))
        return files, labels
